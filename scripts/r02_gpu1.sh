#!/bin/bash
# round-2 GPU call 1: regression tests, parity probes, hub probe, block-order probe, C4 bench + ncu
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $O/gpu1_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/gpu1_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu1_rc.txt
timeout 900 python scripts/r02_probe.py parity > $O/gpu1_parity.log 2>&1; echo "parity rc=$?" | tee -a $O/gpu1_rc.txt
for w in 16 4 1; do GE_HOT_WRITERS=$w timeout 300 python scripts/r02_probe.py zipf > $O/gpu1_zipf_w$w.log 2>&1; echo "zipf$w rc=$?" | tee -a $O/gpu1_rc.txt; done
timeout 900 python scripts/r02_probe.py blocks > $O/gpu1_blocks.log 2>&1; echo "blocks rc=$?" | tee -a $O/gpu1_rc.txt
timeout 600 python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu-baseline > $O/gpu1_bench_c4.json 2> $O/gpu1_bench_c4.log &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:glove_adagrad_update -s 2 -c 1 -o $O/r02_update_c4 \
    python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu-baseline > $O/gpu1_ncu.log 2>&1; echo "ncu rc=$?" | tee -a $O/gpu1_rc.txt
tail -3 $O/gpu1_pytest.log; tail -5 $O/gpu1_parity.log; cat $O/gpu1_bench_c4.json | head -c 1500
