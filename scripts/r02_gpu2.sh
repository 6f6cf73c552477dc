#!/bin/bash
# round-2 GPU call 2 (2 GPUs): copy-engine rotation — hardware correctness test, C4 at N=2 with the 1-rank loss check
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
nvidia-smi topo -m > $O/gpu2_topo.txt 2>&1
GE_DEBUG=1 timeout 900 python -m pytest tests/test_gpu_dist_multi.py -x -q -s > $O/gpu2_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu2_rc.txt
tail -15 $O/gpu2_pytest.log
for rot in ce nccl; do
  GE_DIST_ROT=$rot timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
     bench.py --gpus 2 --steps 4 --warmup 2 --workload c4 > $O/gpu2_bench_c4_2gpu_$rot.json 2> $O/gpu2_bench_c4_2gpu_$rot.log
  echo "bench $rot rc=$?" | tee -a $O/gpu2_rc.txt
  head -c 2500 $O/gpu2_bench_c4_2gpu_$rot.json; echo
done
