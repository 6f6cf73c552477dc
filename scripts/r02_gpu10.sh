#!/bin/bash
# hub defaults (8 on the hottest id; 4 in buckets): probe + the hub test + dist tests, then C4 1-GPU bench
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 500 python scripts/r02_probe.py zipf > $O/gpu10_zipf.log 2>&1; echo "zipf rc=$?" | tee -a $O/gpu10_rc.txt
tail -1 $O/gpu10_zipf.log | python -c "
import sys,json,numpy as np
d=json.loads(sys.stdin.read())
print('hub', d['hub'], 'dist', d['dist_hub'])
for k in ('atomic','dist8x8'):
    print(k, np.round(d[k],5), 'rel', np.round(d[k+'_rel'],4), 'ms', round(d.get(k+'_ms',0),2))
print('dist vs block walk', np.round(d['dist8x8_rel_blockwalk'],4))" | tee -a $O/gpu10_rc.txt
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_dist.py -m gpu -x -q -s > $O/gpu10_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu10_rc.txt
tail -5 $O/gpu10_pytest.log
timeout 900 python bench.py --steps 4 --warmup 3 > $O/gpu10_bench_c4.json 2> $O/gpu10_bench_c4.log; echo "bench c4 rc=$?" | tee -a $O/gpu10_rc.txt
cat $O/gpu10_bench_c4.json
