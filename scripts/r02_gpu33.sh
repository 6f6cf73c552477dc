#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_bca.py -m gpu -x -q > $O/gpu33_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu33_rc.txt
tail -3 $O/gpu33_pytest.log
