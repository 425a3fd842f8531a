#!/bin/sh
# Build libpyccb200.so in-tree for sm_100a (nvcc cross-compiles without a GPU).  The translation units
# are compiled in parallel; each one's ptxas -v report is kept next to it (*.ptxas.log, git-ignored).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-nvcc}
FLAGS="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Xptxas -v"
UNITS="contract contract_f32 contract_tc05 elementwise triples api"
pids=""
for f in $UNITS; do
  ( $NVCC $FLAGS -c $f.cu -o $f.o 2> $f.ptxas.log || { cat $f.ptxas.log; exit 1; } ) &
  pids="$pids $!"
done
for p in $pids; do wait $p; done
OBJS=""
for f in $UNITS; do OBJS="$OBJS $f.o"; done
$NVCC -shared -o ../libpyccb200.so $OBJS
grep -E "registers|spill" contract.ptxas.log | head -20
