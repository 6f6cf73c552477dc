#!/bin/bash
# GPU call 36: ge_flatten_edges_static beside ge_set_graph_edges -> profiles/r02d_static_flatten.md
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 90 python scripts/r02_static_flatten_bench.py > $O/r02d_static_flatten.json 2> $O/r02d_static_flatten.err; echo "rc=$?"
cat $O/r02d_static_flatten.json
