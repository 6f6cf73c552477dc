#!/bin/bash
# development aid: A/B of multi-GPU trainer variants. usage: dist_ab.sh NGPU "ENV=.. ENV=.." workload tag
n=$1; envs=$2; wl=$3; tag=$4
env $envs python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n --steps 4 --warmup 3 --workload $wl > gpurun_out/ab_$tag.json 2> gpurun_out/ab_$tag.err
echo "$tag rc=$?"
