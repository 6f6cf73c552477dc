#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_bca.py tests/test_golden.py -m gpu -x -q > $O/gpu21_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu21_rc.txt
tail -3 $O/gpu21_pytest.log
for t2 in 32 64; do
  GE_BCA_T2_WARPS=$t2 GE_DEBUG=1 timeout 900 python scripts/r02_bca_sweep.py c4 104 232 472 > $O/gpu21_bca_c4_t2w$t2.log 2>&1; echo "bca c4 t2=$t2 rc=$?" | tee -a $O/gpu21_rc.txt
  grep "^\[bca\]\|^[0-9]" $O/gpu21_bca_c4_t2w$t2.log | awk 'NR%3!=1' | cut -c1-250
done
GE_DEBUG=1 timeout 900 python scripts/r02_bca_sweep.py c1 104 232 > $O/gpu21_bca_c1.log 2>&1; echo "bca c1 rc=$?" | tee -a $O/gpu21_rc.txt
grep "^\[bca\]\|^[0-9]" $O/gpu21_bca_c1.log | awk 'NR%3!=1' | cut -c1-250
