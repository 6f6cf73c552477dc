#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_bca.py tests/test_golden.py -m gpu -x -q > $O/gpu20_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu20_rc.txt
tail -3 $O/gpu20_pytest.log
for mid in 0 512 1024 2048; do
  GE_BCA_MID_CAP=$mid GE_DEBUG=1 timeout 900 python scripts/r02_bca_sweep.py c4 104 152 > $O/gpu20_bca_c4_mid$mid.log 2>&1; echo "bca c4 mid=$mid rc=$?" | tee -a $O/gpu20_rc.txt
  grep "^\[bca\]\|^[0-9]" $O/gpu20_bca_c4_mid$mid.log | awk 'NR%3!=1' | cut -c1-250
done
GE_DEBUG=1 timeout 900 python scripts/r02_bca_sweep.py c1 104 > $O/gpu20_bca_c1.log 2>&1; echo "bca c1 rc=$?" | tee -a $O/gpu20_rc.txt
grep "^\[bca\]\|^[0-9]" $O/gpu20_bca_c1.log | awk 'NR%3!=1' | cut -c1-250
