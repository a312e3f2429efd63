#!/usr/bin/env bash
# Marginal cost of each launch class inside the graph-replayed 7B decode step with int8 weights and activations
# (DLI_DEBUG_SKIP drops a class, DLI_A8_SKIP_QUANT the two stand-alone activation quantisers; results garbage; experiments build).
mkdir -p gpurun_out
export DLI_LIB=$(pwd)/distributed_llm_inference_b200/csrc/libdli_b200_exp.so
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --steps 16 --warmup 4 --cpu-budget 0 --no-e2e --a8 ${BENCH_ARGS:-} > gpurun_out/skipa8_$tag.json 2> gpurun_out/skipa8_$tag.err || tail -3 gpurun_out/skipa8_$tag.err
  python -c "
import json
try:
    d=json.loads(open('gpurun_out/skipa8_$tag.json').read().strip().splitlines()[-1]); print('$tag ms/step', d['ms_per_step'])
except Exception as e: print('$tag failed', e)"; }
run none DLI_DEBUG_SKIP=0
run quant DLI_A8_SKIP_QUANT=1
run rmsnorm DLI_DEBUG_SKIP=0x1
run qkv DLI_DEBUG_SKIP=0x2
run attention DLI_DEBUG_SKIP=0x8
run o DLI_DEBUG_SKIP=0x10
run gate_up DLI_DEBUG_SKIP=0x20
run down DLI_DEBUG_SKIP=0x80
run all_gemms DLI_DEBUG_SKIP=0xb2
