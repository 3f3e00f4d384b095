#!/bin/bash
# Build libhybridnet_sm100.so in-tree (sm_100a only).  Usage: csrc/build.sh [-j N]
set -e
cd "$(dirname "$0")"
OUT=../libhybridnet_sm100.so
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --use_fast_math -Xptxas -v"
mkdir -p build
pids=()
for f in hn_*.cu; do
  o=build/${f%.cu}.o
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ] || [ hn_common.cuh -nt "$o" ] || [ hn_attn.cuh -nt "$o" ] || [ ../../include/hybridnet.h -nt "$o" ]; then
    ( $NVCC $FLAGS -c "$f" -o "$o" > "build/${f%.cu}.log" 2>&1 || { cat "build/${f%.cu}.log"; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -o $OUT build/hn_*.o -lcudart
echo "built $OUT"
