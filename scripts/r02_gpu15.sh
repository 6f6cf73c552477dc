#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
for i in 1 2 3; do
  timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -q -s > $O/gpu15_dist_$i.log 2>&1; echo "dist run $i rc=$?" | tee -a $O/gpu15_rc.txt
  grep -h "hub report\|rel vs oracle walking" $O/gpu15_dist_$i.log | cut -c1-300
done
timeout 1500 python -m pytest tests -m gpu -x -q > $O/gpu15_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu15_rc.txt
tail -4 $O/gpu15_pytest.log
timeout 900 python scripts/r02_bca_sweep.py c4 472 232 152 104 > $O/gpu15_bca_c4.log 2>&1; echo "bca c4 rc=$?" | tee -a $O/gpu15_rc.txt
grep -v "^{" $O/gpu15_bca_c4.log | tail -6
timeout 900 python scripts/r02_bca_sweep.py c1 472 232 152 104 > $O/gpu15_bca_c1.log 2>&1; echo "bca c1 rc=$?" | tee -a $O/gpu15_rc.txt
grep -v "^{" $O/gpu15_bca_c1.log | tail -6
