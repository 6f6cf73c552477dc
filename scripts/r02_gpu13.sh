#!/bin/bash
# full GPU suite on the current tree, the random-row microbenchmark, C4 e2e with huge-page advice
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q -s > $O/gpu13_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu13_rc.txt
tail -5 $O/gpu13_pytest.log
grep -n "cyclic\|contiguous\|hub report" $O/gpu13_pytest.log | head
timeout 600 ./scripts/rowbench 10000000 128 400000000 > $O/gpu13_rowbench_d128.json 2> $O/gpu13_rowbench_d128.log; echo "rowbench d128 rc=$?" | tee -a $O/gpu13_rc.txt
cat $O/gpu13_rowbench_d128.json
timeout 600 ./scripts/rowbench 4000000 300 200000000 > $O/gpu13_rowbench_d300.json 2> $O/gpu13_rowbench_d300.log; echo "rowbench d300 rc=$?" | tee -a $O/gpu13_rc.txt
cat $O/gpu13_rowbench_d300.json
timeout 900 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $O/gpu13_bench_c4.json 2> $O/gpu13_bench_c4.log; echo "bench c4 rc=$?" | tee -a $O/gpu13_rc.txt
python -c "
import json; d=json.load(open('$O/gpu13_bench_c4.json')); print(d['ms_per_step'], d['roofline']['frac'], d['e2e'])"
