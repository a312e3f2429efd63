#!/usr/bin/env bash
# Build libdli_b200.so for sm_100a in-tree (nvcc cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v"
mkdir -p build
pids=()
for f in gemm_tcgen05 norm_rope attn_decode attn_decode_gqa attn_prefill attn_prefill_tc span; do
  if [ ! -f build/$f.o ] || [ $f.cu -nt build/$f.o ] || [ common.cuh -nt build/$f.o ] || [ tc_ptx.cuh -nt build/$f.o ] || [ mma_sync.cuh -nt build/$f.o ] || [ ../../include/dli_b200.h -nt build/$f.o ]; then
    ( $NVCC $FLAGS -c $f.cu -o build/$f.o > build/$f.log 2>&1 || { cat build/$f.log; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait $p; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o libdli_b200.so build/*.o
echo "built $(pwd)/libdli_b200.so"
