#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
nproc > $O/gpu24_nproc.txt; free -g | head -2 >> $O/gpu24_nproc.txt
GE_DEBUG=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
   bench.py --gpus 2 --steps 2 --warmup 3 --workload c4 > $O/gpu24_c4_2gpu.json 2> $O/gpu24_c4_2gpu.log; echo "c4 2gpu rc=$?" | tee -a $O/gpu24_rc.txt
grep "^\[dist\|^\[bca\]" $O/gpu24_c4_2gpu.log | head -40
python -c "
import json; d=json.load(open('$O/gpu24_c4_2gpu.json')); print(d['e2e'])"
cat $O/gpu24_nproc.txt
