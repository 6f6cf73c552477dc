#!/bin/bash
# full GPU suite on the current tree, then the ncu launch list of the bench command (after it exited 0 without ncu)
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -x -q > $O/gpu31_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu31_rc.txt
tail -4 $O/gpu31_pytest.log
bash scripts/r02_gpu26.sh
