#!/bin/bash
# 4 GPUs on the shared-memory plumbing: C4 line, N-rank test at world 4, C5-shaped import path
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29553 \
   bench.py --gpus 4 --steps 8 --warmup 3 --workload c4 > $O/gpu32_c4_4gpu.json 2> $O/gpu32_c4_4gpu.log; echo "c4 4gpu rc=$?" | tee -a $O/gpu32_rc.txt
python -c "
import json; d=json.load(open('$O/gpu32_c4_4gpu.json')); print(d['ms_per_step'], round(d['roofline']['frac'],3), d['e2e']['value'], d['e2e']['seconds'], d['e2e']['phases_rank0'], d['config']['check_vs_1rank']['max_rel_loss_diff_vs_1rank'])"
GE_TEST_WORLDS=4 timeout 600 python -m pytest tests/test_gpu_dist_multi.py -x -q -s -k "ce-contiguous" > $O/gpu32_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu32_rc.txt
grep "world=\|passed\|failed" $O/gpu32_pytest.log | cut -c1-330
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29554 \
   bench.py --gpus 4 --steps 3 --warmup 3 --workload c5 > $O/gpu32_c5_4gpu.json 2> $O/gpu32_c5_4gpu.log; echo "c5 4gpu rc=$?" | tee -a $O/gpu32_rc.txt
python -c "
import json; d=json.load(open('$O/gpu32_c5_4gpu.json')); print(d['ms_per_step'], round(d['roofline']['frac'],3), d['e2e']['value'], d['e2e']['seconds'], d['e2e']['phases_rank0'])"
