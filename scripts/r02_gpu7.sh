#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
for cfg in "16 16" "8 8" "4 4" "16 24"; do set -- $cfg
  GE_HOT_WRITERS=$1 GE_HUB_SERIAL=$2 timeout 300 python scripts/r02_probe.py zipf > $O/gpu7_zipf_w$1_hs$2.log 2>&1; echo "zipf w$1 hs$2 rc=$?" | tee -a $O/gpu7_rc.txt
  tail -1 $O/gpu7_zipf_w$1_hs$2.log | python -c "
import sys,json,numpy as np
d=json.loads(sys.stdin.read())
print('atomic', np.round(d['atomic'],5), 'rel', np.round(d['atomic_rel'],4), 'ms', round(d['atomic_ms'],2))"
done
