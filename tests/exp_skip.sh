#!/usr/bin/env bash
# What does each launch class cost inside the graph-replayed 7B decode step?  (DLI_DEBUG_SKIP drops a class; results garbage)
mkdir -p gpurun_out
# environment switches exist only in the experiment build: DLI_EXPERIMENTS=1 distributed_llm_inference_b200/csrc/build.sh
export DLI_LIB=${DLI_LIB:-$(pwd)/distributed_llm_inference_b200/csrc/libdli_b200_exp.so}
for mask in 0 0x1 0x4 0x5 0x8 0x2 0x10 0x20 0x80 0xb2 0xb7 0xd; do
  DLI_DEBUG_SKIP=$mask timeout 300 python bench.py --steps 16 --warmup 4 --cpu-budget 0 ${BENCH_ARGS:-} > gpurun_out/skip_$mask.json 2> gpurun_out/skip_$mask.err || tail -3 gpurun_out/skip_$mask.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/skip_$mask.json")); print("skip $mask ms/step", d["ms_per_step"])
except Exception as e: print("skip $mask failed", e)
PY
done
