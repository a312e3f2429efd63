#!/usr/bin/env bash
# Experiment: int8 activations x int8 weights (kind::i8) next to weight-only int8 and bf16; one box, same clocks.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_i8_gpu.py -q -m gpu -x --no-header -p no:cacheprovider > gpurun_out/i8_tests.log 2>&1; echo "tests exit $?"; tail -n 12 gpurun_out/i8_tests.log
B="--steps 32 --warmup 5 --no-e2e --cpu-budget 0"
for mode in "--w8" "--a8"; do
  name=$(echo $mode | tr -d '-')
  timeout 600 python bench.py $B $mode > gpurun_out/i8_7b_$name.json 2> gpurun_out/i8_7b_$name.err; echo "7b $name $?"
  timeout 600 python bench.py $B $mode --model 70b --layers 10 --batch 32 > gpurun_out/i8_70b_$name.json 2> gpurun_out/i8_70b_$name.err; echo "70b $name $?"
done
python - <<'PY'
import json
for m in ("7b", "70b"):
    for n in ("w8", "a8"):
        try:
            r = json.loads(open(f"gpurun_out/i8_{m}_{n}.json").read().strip().splitlines()[-1])
            print(m, n, r["value"], r["ms_per_step"], r.get("gpu_launches"))
        except Exception as e:
            print(m, n, "failed", e)
PY
