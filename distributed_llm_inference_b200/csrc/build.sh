#!/usr/bin/env bash
# Build libdli_b200.so for sm_100a in-tree (nvcc cross-compiles without a GPU).
# An object is rebuilt when the SHA-256 of (its source, every header, the compiler flags) differs from the one recorded
# beside it -- not by mtime, so objects that travelled with a snapshot are never trusted blindly.
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
if [ "${DLI_EXPERIMENTS:-0}" = "1" ]; then  # A/B and skip switches read from the environment (tests/exp_*.sh); load with DLI_LIB
  FLAGS="$FLAGS -DDLI_EXPERIMENTS"
  OUT=libdli_b200_exp.so
  BUILD=build_exp
fi
if [ -n "${DLI_VARIANT:-}" ]; then  # A/B experiments on one box: -D$DLI_VARIANT into libdli_b200_$DLI_VARIANT.so (load with DLI_LIB)
  FLAGS="$FLAGS -D$DLI_VARIANT"
  OUT=libdli_b200_$DLI_VARIANT.so
  BUILD=build_$DLI_VARIANT
fi
mkdir -p $BUILD
SRCS="gemm_tcgen05 gemm_w8 gemm_i8 norm_rope attn_decode attn_decode_gqa attn_prefill attn_prefill_tc attn_prefill_tc2 span"
HDR_HASH=$(cat *.cuh ../../include/dli_b200.h | sha256sum | cut -d' ' -f1)
pids=()
relink=0
for f in $SRCS; do
  want=$( (echo "$FLAGS $HDR_HASH"; cat $f.cu) | sha256sum | cut -d' ' -f1)
  if [ ! -f $BUILD/$f.o ] || [ "$(cat $BUILD/$f.hash 2>/dev/null)" != "$want" ]; then
    relink=1
    ( $NVCC $FLAGS -c $f.cu -o $BUILD/$f.o > $BUILD/$f.log 2>&1 && echo "$want" > $BUILD/$f.hash || { rm -f $BUILD/$f.hash; cat $BUILD/$f.log; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait $p; done
objs=""
for f in $SRCS; do objs="$objs $BUILD/$f.o"; done
if [ $relink = 1 ] || [ ! -f $OUT ]; then
  $NVCC -shared -gencode arch=compute_100a,code=sm_100a -o $OUT $objs
fi
echo "built $(pwd)/$OUT"
