#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_bca.py tests/test_golden.py -m gpu -x -q > $O/gpu19_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu19_rc.txt
tail -3 $O/gpu19_pytest.log
for t2 in 32 64; do
  GE_BCA_T2_WARPS=$t2 GE_DEBUG=1 timeout 900 python scripts/r02_bca_sweep.py c4 472 232 152 104 > $O/gpu19_bca_c4_t2w$t2.log 2>&1; echo "bca c4 t2=$t2 rc=$?" | tee -a $O/gpu19_rc.txt
  grep "^\[bca\]\|^[0-9]" $O/gpu19_bca_c4_t2w$t2.log | awk 'NR%3!=2' | tail -8
done
GE_BCA_T2_WARPS=32 GE_DEBUG=1 timeout 900 python scripts/r02_bca_sweep.py c1 472 232 152 104 > $O/gpu19_bca_c1.log 2>&1; echo "bca c1 rc=$?" | tee -a $O/gpu19_rc.txt
grep "^\[bca\]\|^[0-9]" $O/gpu19_bca_c1.log | awk 'NR%3!=2' | tail -8
