#!/bin/bash
# full GPU suite + C4 / C3 bench lines on the current defaults (+ smoke)
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -x -q > $O/gpu22_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu22_rc.txt
tail -4 $O/gpu22_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/gpu22_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/gpu22_rc.txt
tail -2 $O/gpu22_smoke.log
timeout 900 python bench.py > $O/gpu22_bench_c4.json 2> $O/gpu22_bench_c4.log; echo "bench c4 rc=$?" | tee -a $O/gpu22_rc.txt
python -c "
import json; d=json.load(open('$O/gpu22_bench_c4.json')); print(d['ms_per_step'], d['roofline']['frac'], d['cooc']['seconds'], d['e2e'], d['cpu_baseline'])"
timeout 900 python bench.py --workload c3 --steps 25 --warmup 3 > $O/gpu22_bench_c3.json 2> $O/gpu22_bench_c3.log; echo "bench c3 rc=$?" | tee -a $O/gpu22_rc.txt
python -c "
import json; d=json.load(open('$O/gpu22_bench_c3.json')); print(d['ms_per_step'], d['roofline']['frac'], d['cooc']['seconds'], d['e2e'])"
timeout 900 python bench.py --impl reference > $O/gpu22_bench_ref.json 2> $O/gpu22_bench_ref.log; echo "bench ref rc=$?" | tee -a $O/gpu22_rc.txt
cut -c1-600 $O/gpu22_bench_ref.json
