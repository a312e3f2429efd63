#!/usr/bin/env bash
# int8 paths after the API change: both quantised test files, then the default one-GPU bench line (with its int8 records)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_i8_gpu.py tests/test_w8_gpu.py -q -m gpu -x --no-header -p no:cacheprovider > gpurun_out/i8w8_tests.log 2>&1; echo "tests exit $?"; tail -n 3 gpurun_out/i8w8_tests.log
timeout 900 python bench.py > gpurun_out/r02v_bench_1gpu.json 2> gpurun_out/r02v_bench_1gpu.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02v_bench_1gpu.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"])
for k in ("decode_int8_weights", "decode_llm_int8"):
    r=d[k]; print(k, r["value"], r["ms_per_step"], r["vs_bf16"], r["launches"])
PY
