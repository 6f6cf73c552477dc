#!/bin/bash
# round-2 8-GPU call: C4 with the copy-engine rotation (1-rank loss check inside), NVLink byte counters
# around it, the NCCL rotation for comparison, C4 at N=4, c5 (configs[4]) at 8 GPUs, the N-rank test.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
nvidia-smi topo -m > $O/gpu11_topo.txt 2>&1
run() { # name nproc extra-env... -- bench args
  local name=$1 n=$2; shift 2
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
     bench.py --gpus $n "$@" > $O/gpu11_$name.json 2> $O/gpu11_$name.log
  echo "$name rc=$?" | tee -a $O/gpu11_rc.txt
  head -c 3000 $O/gpu11_$name.json; echo
}
nvidia-smi nvlink -gt d -i 0 > $O/gpu11_nvlink_before.txt 2>&1
GE_DIST_ROT=ce run c4_8gpu_ce 8 --steps 8 --warmup 3 --workload c4
nvidia-smi nvlink -gt d -i 0 > $O/gpu11_nvlink_after.txt 2>&1
GE_DIST_ROT=nccl run c4_8gpu_nccl 8 --steps 8 --warmup 3 --workload c4
run c5_8gpu 8 --steps 4 --warmup 3 --workload c5
run c4_4gpu_ce 4 --steps 8 --warmup 3 --workload c4
timeout 600 python -m pytest tests/test_gpu_dist_multi.py -x -q -s > $O/gpu11_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu11_rc.txt
tail -6 $O/gpu11_pytest.log
