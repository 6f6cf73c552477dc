#!/bin/bash
# cap criterion = damped share of the launch: Zipf probe, hub + dist tests, hub report of C4
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 500 python scripts/r02_probe.py zipf > $O/gpu14_zipf.log 2>&1; echo "zipf rc=$?" | tee -a $O/gpu14_rc.txt
tail -1 $O/gpu14_zipf.log | python -c "
import sys,json,numpy as np
d=json.loads(sys.stdin.read())
print('hub', d['hub'], 'dist', d['dist_hub'])
for k in ('atomic','dist8x8'):
    print(k, np.round(d[k],5), 'rel', np.round(d[k+'_rel'],4), 'ms', round(d.get(k+'_ms',0),2))
print('dist vs block walk', np.round(d['dist8x8_rel_blockwalk'],4))" | tee -a $O/gpu14_rc.txt
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_dist.py -m gpu -x -q -s > $O/gpu14_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu14_rc.txt
tail -3 $O/gpu14_pytest.log
timeout 600 python scripts/r02_hubinfo.py c4 > $O/gpu14_hubinfo_c4.json 2> $O/gpu14_hubinfo_c4.log; echo "hubinfo rc=$?" | tee -a $O/gpu14_rc.txt
cat $O/gpu14_hubinfo_c4.json
