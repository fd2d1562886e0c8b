#!/bin/bash
# Build libfa_b200.so in-tree for sm_100a (one nvcc per translation unit, in parallel).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -I../include -Icsrc ${FA_NVCC_EXTRA}"
BUILD=${FA_BUILD_DIR:-build}
OUT=${FA_OUT:-vllm_flash_attn/libfa_b200.so}
mkdir -p $BUILD
pids=()
for f in fa_api fa_combine fa_append fa_decode fa_tile fa_prefill_tc5 fa_exchange; do
  if [ ! -f $BUILD/$f.o ] || [ -n "$(find csrc ../include -newer $BUILD/$f.o -type f | head -1)" ]; then
    ( $NVCC $FLAGS -Xptxas -v -c csrc/$f.cu -o $BUILD/$f.o > $BUILD/$f.log 2>&1 || { echo "FAILED $f"; tail -30 $BUILD/$f.log; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o $OUT $BUILD/*.o -lcudart
echo "built $OUT"
