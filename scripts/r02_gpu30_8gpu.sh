#!/bin/bash
# 8 GPUs on the shared-memory plumbing: C4 line with e2e phases, N-rank tests (copy-engine plane, both layouts)
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
GE_DEBUG=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 \
   bench.py --gpus 8 --steps 8 --warmup 3 --workload c4 > $O/gpu30_c4_8gpu.json 2> $O/gpu30_c4_8gpu.log; echo "c4 8gpu rc=$?" | tee -a $O/gpu30_rc.txt
grep "ncclCommInit\|falls back\|^\[dist hand" $O/gpu30_c4_8gpu.log | sort | uniq -c | sort -rn | head -8
python -c "
import json; d=json.load(open('$O/gpu30_c4_8gpu.json')); print(d['ms_per_step'], round(d['roofline']['frac'],3), d['cooc']['seconds'], d['e2e']['value'], d['e2e']['seconds'], d['e2e']['phases_rank0'], d['config']['check_vs_1rank']['max_rel_loss_diff_vs_1rank'], d['config']['parallelism'])"
GE_TEST_WORLDS=8 timeout 600 python -m pytest tests/test_gpu_dist_multi.py -x -q -s -k "ce" > $O/gpu30_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu30_rc.txt
grep "world=\|passed\|failed" $O/gpu30_pytest.log | cut -c1-330
