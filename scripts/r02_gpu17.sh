#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
for t2 in 16 28; do
  GE_BCA_T2_WARPS=$t2 GE_DEBUG=1 timeout 900 python scripts/r02_bca_sweep.py c4 472 312 232 > $O/gpu17_bca_c4_t2w$t2.log 2>&1; echo "bca c4 t2=$t2 rc=$?" | tee -a $O/gpu17_rc.txt
  grep "^\[bca\]\|^[0-9]" $O/gpu17_bca_c4_t2w$t2.log | tail -9
done
