#!/bin/bash
# Build libb200ir.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --use_fast_math=false"
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xptxas -v"
mkdir -p build
OBJS=""
for f in *.cu; do
  o=build/${f%.cu}.o
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ] || [ common.cuh -nt "$o" ] || [ ../../include/b200ir.h -nt "$o" ] || [ tc_common.cuh -nt "$o" ] || [ dwconv_tma.cuh -nt "$o" ]; then
    echo "nvcc $f"
    $NVCC $FLAGS -c "$f" -o "$o" 2> build/${f%.cu}.ptxas.log || { cat build/${f%.cu}.ptxas.log; exit 1; }
  fi
  OBJS="$OBJS $o"
done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o ../libb200ir.so $OBJS
echo "built $(cd ..; pwd)/libb200ir.so"
