#!/usr/bin/env bash
mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?"; tail -n ${TAILN:-12} gpurun_out/$name.log | cut -c1-${CUTC:-400}; }
run ops python -m pytest tests/test_ops_gpu.py -q -m gpu --no-header -p no:cacheprovider --maxfail=10
run span python -m pytest tests/test_span_gpu.py tests/test_zz_no_alloc.py -q -m gpu --no-header -p no:cacheprovider -s --maxfail=10
for b in 16 1; do
  timeout 600 python bench.py --steps 32 --warmup 4 --cpu-budget 2 --batch $b > gpurun_out/bench_b$b.json 2> gpurun_out/bench_b$b.err || tail -5 gpurun_out/bench_b$b.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_b$b.json"))
    print("B=$b value",d["value"],"ms/step",d["ms_per_step"],"stage",d["stage_roofline"]["frac"],"gemm",d["roofline"]["frac"],d["roofline"]["ms_per_launch"],"e2e",d["e2e"]["value"],"prefill",d["prefill_tokens_per_sec"])
    print("   ",d["kernel_ms_per_step"])
except Exception as e: print("bench B=$b failed",e)
PY
done
