#!/bin/bash
# final 8-GPU state: C4 and C5 lines with e2e phases, the N-rank test at world 8 (ce / nccl / cyclic)
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
nproc > $O/gpu25_host.txt; free -g | head -2 >> $O/gpu25_host.txt
run() { local name=$1 n=$2; shift 2
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29537 \
     bench.py --gpus $n "$@" > $O/gpu25_$name.json 2> $O/gpu25_$name.log
  echo "$name rc=$?" | tee -a $O/gpu25_rc.txt
  python -c "
import json; d=json.load(open('$O/gpu25_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],3), d['cooc'] and d['cooc']['seconds'], d['e2e']['value'], d['e2e']['seconds'], d['e2e']['phases_rank0'], d['config']['check_vs_1rank'] and d['config']['check_vs_1rank']['max_rel_loss_diff_vs_1rank'], d['config']['hub'])"
}
run c4_8gpu 8 --steps 8 --warmup 3 --workload c4
run c5_8gpu 8 --steps 4 --warmup 3 --workload c5
GE_TEST_WORLDS=8 timeout 600 python -m pytest tests/test_gpu_dist_multi.py -x -q -s > $O/gpu25_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu25_rc.txt
grep "world=\|passed\|failed" $O/gpu25_pytest.log | cut -c1-330
cat $O/gpu25_host.txt
