#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
for i in 1 2 3; do
  timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -q -s -k block_schedule > $O/gpu16_dist_$i.log 2>&1; echo "dist run $i rc=$?" | tee -a $O/gpu16_rc.txt
  grep -h "rel vs oracle walking" $O/gpu16_dist_$i.log | cut -c1-300
done
GE_DEBUG=1 timeout 900 python scripts/r02_bca_sweep.py c4 472 640 960 > $O/gpu16_bca_c4.log 2>&1; echo "bca c4 rc=$?" | tee -a $O/gpu16_rc.txt
grep "^\[bca\]\|^[0-9]" $O/gpu16_bca_c4.log | tail -12
