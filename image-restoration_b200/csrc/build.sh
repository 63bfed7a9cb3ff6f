#!/bin/bash
# Build libb200ir.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --use_fast_math=false"
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xptxas -v $EXTRA_FLAGS"
# A/B builds: EXTRA_FLAGS="-D..." BUILD=build_x OUT=../../tools/ab/libx.so bash build.sh ; run with B200IR_LIB=tools/ab/libx.so
BUILD=${BUILD:-build}
OUT=${OUT:-../libb200ir.so}
mkdir -p $BUILD "$(dirname $OUT)"
OBJS=""
PIDS=""
for f in *.cu; do
  o=$BUILD/${f%.cu}.o
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ] || [ common.cuh -nt "$o" ] || [ ../../include/b200ir.h -nt "$o" ] || [ tc_common.cuh -nt "$o" ] || [ dwconv_tma.cuh -nt "$o" ]; then
    echo "nvcc $f"
    ( $NVCC $FLAGS -c "$f" -o "$o.tmp.o" 2> $BUILD/${f%.cu}.ptxas.log && mv "$o.tmp.o" "$o" || { cat $BUILD/${f%.cu}.ptxas.log; rm -f "$o" "$o.tmp.o"; exit 1; } ) &
    PIDS="$PIDS $!"
  fi
  OBJS="$OBJS $o"
done
for p in $PIDS; do wait $p || exit 1; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o $OUT $OBJS
echo "built $OUT"
