#!/bin/bash
# round-2 GPU call 5 (1 GPU): full GPU test-suite, L2 keep sweep with hints on hot rows only, C4 / C1 bench lines
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -s > $O/gpu5_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu5_rc.txt
tail -5 $O/gpu5_pytest.log
PROBE_KEEP=0,32,64 timeout 600 python scripts/r02_probe.py keep > $O/gpu5_keep_c4.log 2>&1; echo "keep c4 rc=$?" | tee -a $O/gpu5_rc.txt
cat $O/gpu5_keep_c4.log
timeout 600 python bench.py --workload c4 --steps 4 --warmup 3 > $O/gpu5_bench_c4.json 2> $O/gpu5_bench_c4.log; echo "bench c4 rc=$?" | tee -a $O/gpu5_rc.txt
head -c 2600 $O/gpu5_bench_c4.json; echo
timeout 900 python bench.py --workload c1 --steps 8 --warmup 3 > $O/gpu5_bench_c1_d300.json 2> $O/gpu5_bench_c1_d300.log; echo "bench c1 rc=$?" | tee -a $O/gpu5_rc.txt
head -c 1500 $O/gpu5_bench_c1_d300.json; echo
