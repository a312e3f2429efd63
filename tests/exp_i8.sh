#!/usr/bin/env bash
# Experiment: int8 activations x int8 weights (kind::i8) next to weight-only int8; one box, same clocks.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_i8_gpu.py -q -m gpu -x --no-header -p no:cacheprovider > gpurun_out/i8_tests.log 2>&1; echo "tests exit $?"; tail -n 4 gpurun_out/i8_tests.log
B="--steps 32 --warmup 5 --no-e2e --cpu-budget 0"
for mode in "--w8" "--a8" "--a8 --a8-threshold 0"; do
  name=$(echo $mode | tr -d ' -')
  timeout 600 python bench.py $B $mode > gpurun_out/i8_7b_$name.json 2> gpurun_out/i8_7b_$name.err; echo "7b $name $?"
  timeout 600 python bench.py $B $mode --model 70b --layers 10 --batch 32 > gpurun_out/i8_70b_$name.json 2> gpurun_out/i8_70b_$name.err; echo "70b $name $?"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/i8_7*b_*.json")):
    try:
        r = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, r["value"], r["ms_per_step"], r.get("gpu_launches"))
    except Exception as e:
        print(f, "failed", e)
PY
timeout 300 python tests/exp_gemm_w8.py > gpurun_out/i8_gemm_micro.log 2>&1; tail -n 5 gpurun_out/i8_gemm_micro.log | cut -c1-260
if [ "${I8_NCU:-0}" = "1" ]; then
  python tests/prof_step.py decode_a8 > gpurun_out/prof_plain_decode_a8.log 2>&1 || { tail -5 gpurun_out/prof_plain_decode_a8.log; exit 1; }
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
      --log-file gpurun_out/launches_decode_a8.csv python tests/prof_step.py decode_a8 > gpurun_out/ncu_decode_a8.log 2>&1; echo "launch list a8: $?"
  ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_i8_kernel -c 4 \
      -o gpurun_out/prof_gemm_i8 -f python tests/prof_step.py decode_a8 > gpurun_out/ncu_full_i8.log 2>&1; echo "full i8: $?"
fi
