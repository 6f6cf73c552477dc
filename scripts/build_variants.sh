#!/bin/bash
# development aid: libgraphemb_o<N>.so with -DGE_LOAD_ORDER=N (issue order of the row loads in K6)
set -e
cd "$(dirname "$0")/../graph_embeddings_b200/csrc"
for n in "$@"; do
  d=$(mktemp -d)
  for f in api train bca dist writer graph; do
    /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -DGE_LOAD_ORDER=$n -c $f.cu -o $d/$f.o &
  done
  wait
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libgraphemb_o$n.so $d/*.o -lcudart
  rm -rf $d
done
