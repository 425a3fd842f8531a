#!/bin/sh
# Build libpyccb200.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-nvcc}
FLAGS="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Xptxas -v"
for f in contract contract_f32 contract_tc05 elementwise triples api; do
  $NVCC $FLAGS -c $f.cu -o $f.o 2> $f.ptxas.log || { cat $f.ptxas.log; exit 1; }
done
$NVCC -shared -o ../libpyccb200.so contract.o contract_f32.o contract_tc05.o elementwise.o triples.o api.o
grep -E "registers|spill" contract.ptxas.log | head -20
