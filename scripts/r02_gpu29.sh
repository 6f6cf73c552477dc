#!/bin/bash
# writers = GE_HUB_LOAD / share: Zipf probe, hub test + dist tests, C3 8x8 on one GPU (capped buckets, ms/epoch)
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 600 python scripts/r02_probe.py zipf > $O/gpu29_zipf.log 2>&1; echo "zipf rc=$?" | tee -a $O/gpu29_rc.txt
tail -1 $O/gpu29_zipf.log | python -c "
import sys,json,numpy as np
d=json.loads(sys.stdin.read())
print('hub', d['hub'], 'dist', d['dist_hub'])
for k in ('atomic','dist8x8'):
    print(k, np.round(d[k],5), 'rel', np.round(d[k+'_rel'],4), 'ms', round(d.get(k+'_ms',0),2))
print('dist vs block walk', np.round(d['dist8x8_rel_blockwalk'],4))" | tee -a $O/gpu29_rc.txt
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_dist.py -m gpu -x -q > $O/gpu29_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu29_rc.txt
tail -3 $O/gpu29_pytest.log
PROBE_WL=c3 timeout 600 python scripts/r02_probe.py blocks > $O/gpu29_blocks_c3.log 2>&1; echo "blocks c3 rc=$?" | tee -a $O/gpu29_rc.txt
tail -1 $O/gpu29_blocks_c3.log | cut -c1-900
timeout 300 python scripts/r02_hubinfo.py c3 2>/dev/null | tail -1
