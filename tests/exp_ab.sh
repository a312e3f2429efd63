#!/usr/bin/env bash
# A/B two builds of the library on ONE box (boxes differ by several percent in tensor-bound work): alternates them.
# usage: exp_ab.sh <alt lib path> [bench args...]
ALT=$1; shift
mkdir -p gpurun_out
for rep in 1 2; do
  for which in main alt; do
    if [ $which = alt ]; then export DLI_LIB=$ALT; else unset DLI_LIB; fi
    timeout 600 python bench.py --steps 8 --warmup 3 --cpu-budget 0 "$@" > gpurun_out/ab_$which.json 2> gpurun_out/ab_$which.err || tail -3 gpurun_out/ab_$which.err
    python - <<PY
import json
d=json.load(open("gpurun_out/ab_$which.json"))
p=d["prefill"]
print("$which rep $rep: decode ms/step", d["ms_per_step"], "prefill ms", p["ms"], {k: v for k, v in p["kernel_ms"].items() if v > 0.5})
PY
  done
done
