#!/bin/bash
# warp cap (onset 64 -> 16 / 8 / 24 on the hottest id) on the Zipf(1) matrix: single launch and 8x8 schedule vs the oracle
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
for cfg in "16 16" "16 8" "16 24"; do set -- $cfg
  GE_HOT_WRITERS=$1 GE_HUB_SERIAL=$2 timeout 400 python scripts/r02_probe.py zipf > $O/gpu9_zipf_w$1_hs$2.log 2>&1; echo "zipf w$1 hs$2 rc=$?" | tee -a $O/gpu9_rc.txt
  tail -1 $O/gpu9_zipf_w$1_hs$2.log | python -c "
import sys,json,numpy as np
d=json.loads(sys.stdin.read())
print('hub', d['hub'], 'dist', d['dist_hub'])
for k in ('atomic','dist8x8'):
    print(k, np.round(d[k],5), 'rel', np.round(d[k+'_rel'],4), 'ms', round(d.get(k+'_ms',0),2))
print('dist vs block walk', np.round(d['dist8x8_rel_blockwalk'],4))" | tee -a $O/gpu9_rc.txt
done
