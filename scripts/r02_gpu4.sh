#!/bin/bash
# round-2 GPU call 4 (1 GPU): full GPU test-suite (BCA active list, staged hand-off, parity at scale), hub probe with the
# state-dependent stale-step correction, L2 keep sweep with hints on hot rows only, default C4 bench line
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -s > $O/gpu4_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu4_rc.txt
tail -5 $O/gpu4_pytest.log
for w in 16 64; do GE_HOT_WRITERS=$w timeout 300 python scripts/r02_probe.py zipf > $O/gpu4_zipf_w$w.log 2>&1; echo "zipf$w rc=$?" | tee -a $O/gpu4_rc.txt; done
PROBE_KEEP=0,64 timeout 600 python scripts/r02_probe.py keep > $O/gpu4_keep_c4.log 2>&1; echo "keep c4 rc=$?" | tee -a $O/gpu4_rc.txt
cat $O/gpu4_keep_c4.log
timeout 600 python bench.py --workload c4 --steps 3 --warmup 3 > $O/gpu4_bench_c4.json 2> $O/gpu4_bench_c4.log; echo "bench rc=$?" | tee -a $O/gpu4_rc.txt
head -c 2500 $O/gpu4_bench_c4.json
