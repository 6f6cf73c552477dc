#!/bin/bash
# GPU call 34: the allocation-free entry point alone in a fresh process, then the whole -m gpu suite (it runs last)
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 170 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/zz_alone.log 2>&1; echo "rc=$?" >> $O/zz_alone.log
tail -3 $O/zz_alone.log
timeout 280 python -m pytest tests -m gpu -x -q --durations=8 > $O/zz_full.log 2>&1; echo "rc=$?" >> $O/zz_full.log
tail -15 $O/zz_full.log
