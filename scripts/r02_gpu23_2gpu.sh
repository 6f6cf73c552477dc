#!/bin/bash
# 2 GPUs: C4 e2e phases, the N-rank tests (contiguous / nccl / cyclic); on GPU 0: Adam/AMSGrad Hogwild spread (3 runs)
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
   bench.py --gpus 2 --steps 8 --warmup 3 --workload c4 > $O/gpu23_c4_2gpu.json 2> $O/gpu23_c4_2gpu.log; echo "c4 2gpu rc=$?" | tee -a $O/gpu23_rc.txt
python -c "
import json; d=json.load(open('$O/gpu23_c4_2gpu.json')); print(d['ms_per_step'], d['roofline']['frac'], d['cooc'], d['e2e'], d['config']['check_vs_1rank'])"
timeout 900 python -m pytest tests/test_gpu_dist_multi.py -x -q -s > $O/gpu23_pytest_multi.log 2>&1; echo "multi rc=$?" | tee -a $O/gpu23_rc.txt
grep "world=\|passed\|failed" $O/gpu23_pytest_multi.log | cut -c1-330
for i in 1 2 3; do
  timeout 900 python -m pytest tests/test_gpu_optimizers.py -q -s -k hogwild > $O/gpu23_opt_$i.log 2>&1; echo "opt run $i rc=$?" | tee -a $O/gpu23_rc.txt
  grep "rel gpu" $O/gpu23_opt_$i.log | cut -c1-400
done
