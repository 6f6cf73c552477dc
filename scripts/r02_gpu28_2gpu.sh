#!/bin/bash
# 2 GPUs: node plumbing through shared memory (no NCCL communicator on the copy-engine plane): N-rank tests, C4 phases
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
GE_DEBUG=1 timeout 900 python -m pytest tests/test_gpu_dist_multi.py -x -q -s > $O/gpu28_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu28_rc.txt
grep "world=\|passed\|failed\|ncclCommInit\|falls back" $O/gpu28_pytest.log | cut -c1-300 | head -20
tail -5 $O/gpu28_pytest.log | cut -c1-300
GE_DEBUG=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 \
   bench.py --gpus 2 --steps 8 --warmup 3 --workload c4 > $O/gpu28_c4_2gpu.json 2> $O/gpu28_c4_2gpu.log; echo "c4 2gpu rc=$?" | tee -a $O/gpu28_rc.txt
grep "^\[dist\|ncclCommInit\|falls back" $O/gpu28_c4_2gpu.log | sort | uniq -c | sort -rn | head -14
python -c "
import json; d=json.load(open('$O/gpu28_c4_2gpu.json')); print(d['ms_per_step'], round(d['roofline']['frac'],3), d['e2e']['value'], d['e2e']['seconds'], d['e2e']['phases_rank0'], d['config']['check_vs_1rank']['max_rel_loss_diff_vs_1rank'], d['config']['parallelism'])"
GE_DIST_PLUMBING=nccl timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 \
   bench.py --gpus 2 --steps 2 --warmup 3 --workload c3 > $O/gpu28_c3_2gpu_ncclplumb.json 2> $O/gpu28_c3_2gpu_ncclplumb.log; echo "c3 nccl plumbing rc=$?" | tee -a $O/gpu28_rc.txt
python -c "
import json; d=json.load(open('$O/gpu28_c3_2gpu_ncclplumb.json')); print(d['ms_per_step'], d['e2e']['phases_rank0'], d['config']['check_vs_1rank']['max_rel_loss_diff_vs_1rank'])"
