#!/bin/bash
# Build variants of one kernel file (compile-time experiment macros; default hn_attn_tc.cu, SRC=hn_conv_tc.cu for another)
# as separate shared objects under scripts/build/ for scripts/attn_bench.py --lib / scripts/conv_bench.py --lib.
# Usage: [SRC=file.cu] scripts/attn_variants.sh name "-DX=1 -DY=2" ...
set -e
cd "$(dirname "$0")/../transformer-cnn-hybrid-network-for-video-processing_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --use_fast_math"
OUT=../../scripts/build
mkdir -p $OUT
SRC=${SRC:-hn_attn_tc.cu}
OBJS=$(ls build/hn_*.o | grep -v ${SRC%.cu}.o)
while [ $# -ge 2 ]; do
  name=$1; defs=$2; shift 2
  ( $NVCC $FLAGS $defs -c $SRC -o $OUT/var_$name.o && $NVCC -shared -o $OUT/libvar_$name.so $OBJS $OUT/var_$name.o -lcudart && echo "built $name ($defs)" ) &
done
wait
