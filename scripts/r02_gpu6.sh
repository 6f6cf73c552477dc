#!/bin/bash
# round-2 GPU call 6 (1 GPU): extreme-hub segment (bounded concurrency) on the Zipf(1) matrix, full GPU tests
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
for hs in 64 32 0; do GE_HUB_SERIAL=$hs timeout 300 python scripts/r02_probe.py zipf > $O/gpu6_zipf_hs$hs.log 2>&1; echo "zipf hs$hs rc=$?" | tee -a $O/gpu6_rc.txt; tail -1 $O/gpu6_zipf_hs$hs.log | cut -c1-1500; done
timeout 1500 python -m pytest tests -m gpu -q -s > $O/gpu6_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu6_rc.txt
tail -5 $O/gpu6_pytest.log
