#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
for t2 in 32 48 64; do
  GE_BCA_T2_WARPS=$t2 GE_DEBUG=1 timeout 900 python scripts/r02_bca_sweep.py c4 232 152 > $O/gpu18_bca_c4_t2w$t2.log 2>&1; echo "bca c4 t2=$t2 rc=$?" | tee -a $O/gpu18_rc.txt
  grep "^\[bca\]\|^[0-9]" $O/gpu18_bca_c4_t2w$t2.log | tail -6
done
GE_BCA_T2_WARPS=48 GE_DEBUG=1 timeout 900 python scripts/r02_bca_sweep.py c1 472 232 152 > $O/gpu18_bca_c1.log 2>&1; echo "bca c1 rc=$?" | tee -a $O/gpu18_rc.txt
grep "^\[bca\]\|^[0-9]" $O/gpu18_bca_c1.log | tail -9
