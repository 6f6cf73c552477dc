#!/bin/bash
# GPU call 37: drain threads of the f64 extract (8 vs 16) on a C4-sized store, parity-scale test through the
# threaded drain, default bench line on the same box
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 100 python scripts/r02_extract_threads.py > $O/r02e_extract_threads.json 2> $O/r02e_extract_threads.err; echo "rc=$?"
cat $O/r02e_extract_threads.json; tail -2 $O/r02e_extract_threads.err
timeout 100 python bench.py > $O/r02e_bench_c4.json 2> $O/r02e_bench.err; echo "rc=$?"
python -c "
import json;d=json.loads(open('$O/r02e_bench_c4.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['roofline']['frac'],d['e2e'])"
timeout 60 python -m pytest tests/test_gpu_parity_scale.py -m gpu -x -q -k c1_pglove_d300 > $O/r02e_parity.log 2>&1; echo "rc=$?"; tail -2 $O/r02e_parity.log
