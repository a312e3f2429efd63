#!/usr/bin/env bash
# Build libdli_b200.so for sm_100a in-tree (nvcc cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v"
OUT=libdli_b200.so
BUILD=build
if [ "${DLI_TRACE:-0}" = "1" ]; then  # timeline build for tests/exp_trace.py; never the product library
  FLAGS="$FLAGS -DDLI_TRACE"
  OUT=libdli_b200_trace.so
  BUILD=build_trace
fi
if [ -n "${DLI_VARIANT:-}" ]; then  # A/B experiments on one box: -D$DLI_VARIANT into libdli_b200_$DLI_VARIANT.so (load with DLI_LIB)
  FLAGS="$FLAGS -D$DLI_VARIANT"
  OUT=libdli_b200_$DLI_VARIANT.so
  BUILD=build_$DLI_VARIANT
fi
mkdir -p $BUILD
pids=()
for f in gemm_tcgen05 norm_rope attn_decode attn_decode_gqa attn_prefill attn_prefill_tc attn_prefill_tc2 span; do
  if [ ! -f $BUILD/$f.o ] || [ $f.cu -nt $BUILD/$f.o ] || [ common.cuh -nt $BUILD/$f.o ] || [ tc_ptx.cuh -nt $BUILD/$f.o ] || [ mma_sync.cuh -nt $BUILD/$f.o ] || [ rope_inline.cuh -nt $BUILD/$f.o ] || [ ../../include/dli_b200.h -nt $BUILD/$f.o ]; then
    ( $NVCC $FLAGS -c $f.cu -o $BUILD/$f.o > $BUILD/$f.log 2>&1 || { cat $BUILD/$f.log; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait $p; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o $OUT $BUILD/*.o
echo "built $(pwd)/$OUT"
