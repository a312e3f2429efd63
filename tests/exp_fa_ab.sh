#!/usr/bin/env bash
# same-box A/B of the two tcgen05 prefill attention kernels (DLI_FA_V1=1 selects the first generation)
mkdir -p gpurun_out
# environment switches exist only in the experiment build: DLI_EXPERIMENTS=1 distributed_llm_inference_b200/csrc/build.sh
export DLI_LIB=${DLI_LIB:-$(pwd)/distributed_llm_inference_b200/csrc/libdli_b200_exp.so}
for rep in 1 2; do
  for v1 in "" 1; do
    if [ -n "$v1" ]; then export DLI_FA_V1=1; else unset DLI_FA_V1; fi
    for cfg in "--model 7b" "--model 70b --layers 10 --batch 1 --prompt 8192"; do
      timeout 600 python bench.py --steps 4 --warmup 3 --cpu-budget 0 $cfg > gpurun_out/fa_ab.json 2> gpurun_out/fa_ab.err || tail -3 gpurun_out/fa_ab.err
      python - <<PY
import json
d=json.load(open("gpurun_out/fa_ab.json")); p=d["prefill"]
print("v1='$v1' $cfg: prefill ms", p["ms"], "attention", p["kernel_ms"]["attention"], "gate_up", p["kernel_ms"]["gemm_gate_up"])
PY
    done
  done
done
