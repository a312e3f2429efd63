#!/usr/bin/env bash
# Experiment: operand-ring size of the int8 x int8 GEMM (how many CTAs of consecutive launches can be co-resident per SM).
mkdir -p gpurun_out
E=$PWD/distributed_llm_inference_b200/csrc/libdli_b200_exp.so
timeout 600 python -m pytest tests/test_i8_gpu.py -q -m gpu -x --no-header -p no:cacheprovider -k "quantize_act or span" > gpurun_out/i8_tests2.log 2>&1; echo "tests exit $?"; tail -n 3 gpurun_out/i8_tests2.log
B="--steps 32 --warmup 5 --no-e2e --cpu-budget 0 --a8"
for kb in ${RINGS:-110 72 54 40}; do
  DLI_LIB=$E DLI_I8_RING_KB=$kb timeout 600 python bench.py $B > gpurun_out/i8ring_7b_$kb.json 2> gpurun_out/i8ring_7b_$kb.err; echo "7b ring $kb: $?"
  DLI_LIB=$E DLI_I8_RING_KB=$kb timeout 600 python bench.py $B --model 70b --layers 10 --batch 32 > gpurun_out/i8ring_70b_$kb.json 2> gpurun_out/i8ring_70b_$kb.err; echo "70b ring $kb: $?"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/i8ring_*.json")):
    try:
        r = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, r["value"], r["ms_per_step"])
    except Exception as e:
        print(f, "failed", e)
PY
