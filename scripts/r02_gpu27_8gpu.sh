#!/bin/bash
# 8 GPUs: communicator set-up without NVLS (e2e phases), then the N-rank test at world 8
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
GE_DEBUG=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \
   bench.py --gpus 8 --steps 8 --warmup 3 --workload c4 > $O/gpu27_c4_8gpu.json 2> $O/gpu27_c4_8gpu.log; echo "c4 8gpu rc=$?" | tee -a $O/gpu27_rc.txt
grep "^\[dist init\|^\[dist hand" $O/gpu27_c4_8gpu.log | sort | uniq -c | sort -rn | head -12
python -c "
import json; d=json.load(open('$O/gpu27_c4_8gpu.json')); print(d['ms_per_step'], round(d['roofline']['frac'],3), d['e2e']['value'], d['e2e']['seconds'], d['e2e']['phases_rank0'], d['config']['check_vs_1rank']['max_rel_loss_diff_vs_1rank'])"
GE_TEST_WORLDS=8 timeout 600 python -m pytest tests/test_gpu_dist_multi.py -x -q -s > $O/gpu27_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu27_rc.txt
grep "world=\|passed\|failed" $O/gpu27_pytest.log | cut -c1-330
