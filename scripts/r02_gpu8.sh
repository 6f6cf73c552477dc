#!/bin/bash
# warp-cap sweep on the Zipf(1) hub matrix (single launch, capped concurrency), then the hub + dist tests
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
for cfg in "16 64" "16 32" "16 16" "16 0"; do set -- $cfg
  GE_HOT_WRITERS=$1 GE_HUB_SERIAL=$2 timeout 300 python scripts/r02_probe.py zipf > $O/gpu8_zipf_w$1_hs$2.log 2>&1; echo "zipf w$1 hs$2 rc=$?" | tee -a $O/gpu8_rc.txt
  tail -1 $O/gpu8_zipf_w$1_hs$2.log | python -c "
import sys,json,numpy as np
d=json.loads(sys.stdin.read())
for k in ('atomic','dist8x8'):
    print(k, np.round(d[k],5), 'rel', np.round(d[k+'_rel'],4), 'ms', round(d.get(k+'_ms',0),2))" | tee -a $O/gpu8_rc.txt
done
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_dist.py -m gpu -x -q -s > $O/gpu8_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu8_rc.txt
tail -5 $O/gpu8_pytest.log
