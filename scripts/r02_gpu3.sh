#!/bin/bash
# round-2 GPU call 3 (1 GPU): L2-resident hot rows sweep (C4, C3), hub probe against the exact oracle loss, c1/c2 bench lines
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 600 python scripts/r02_probe.py keep > $O/gpu3_keep_c4.log 2>&1; echo "keep c4 rc=$?" | tee -a $O/gpu3_rc.txt
PROBE_WL=c3 timeout 300 python scripts/r02_probe.py keep > $O/gpu3_keep_c3.log 2>&1; echo "keep c3 rc=$?" | tee -a $O/gpu3_rc.txt
for w in 16 4; do GE_HOT_WRITERS=$w timeout 300 python scripts/r02_probe.py zipf > $O/gpu3_zipf_w$w.log 2>&1; echo "zipf$w rc=$?" | tee -a $O/gpu3_rc.txt; done
timeout 600 python bench.py --workload c2 --warmup 1 > $O/gpu3_bench_c2.json 2> $O/gpu3_bench_c2.log; echo "c2 rc=$?" | tee -a $O/gpu3_rc.txt
timeout 900 python bench.py --workload c1 --dim 100 --steps 8 --warmup 3 --cpu-1thread > $O/gpu3_bench_c1_d100.json 2> $O/gpu3_bench_c1_d100.log; echo "c1 d100 rc=$?" | tee -a $O/gpu3_rc.txt
timeout 600 python -m pytest tests/test_gpu_parity_scale.py tests/test_gpu_train.py -q -s > $O/gpu3_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu3_rc.txt
cat $O/gpu3_keep_c4.log $O/gpu3_keep_c3.log; tail -3 $O/gpu3_pytest.log; head -c 3000 $O/gpu3_bench_c1_d100.json
