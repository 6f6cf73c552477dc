#!/bin/bash
# ncu launch list of the bench command (C4, after the same command exited 0 without ncu) + ncu --set full of the builder's global tier at C3
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/gpu26_bench.json 2> $O/gpu26_bench.log; rc=$?; echo "bench rc=$rc" | tee -a $O/gpu26_rc.txt
if [ $rc -eq 0 ]; then
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/gpu26_launches_c4.csv \
     python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/gpu26_ncu.log 2>&1; echo "ncu launches rc=$?" | tee -a $O/gpu26_rc.txt
  python scripts/summarize_ncu.py launches $O/gpu26_launches_c4.csv $O/gpu26_launches_c4.md; head -16 $O/gpu26_launches_c4.md
fi
