#!/bin/bash
# GPU call 35: smoke() and the default bench line on the round's final build -> profiles/r02d_bench_c4.json
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02d_smoke.log 2>&1; echo "rc=$?" >> $O/r02d_smoke.log
tail -2 $O/r02d_smoke.log
timeout 200 python bench.py > $O/r02d_bench_c4.json 2> $O/r02d_bench.err; echo "rc=$?"
