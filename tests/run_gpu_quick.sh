#!/usr/bin/env bash
# quick loop: decode-attention parity + the three bench configs (no prefill detail)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_zz_no_alloc.py -q -m gpu --no-header -p no:cacheprovider -k "${K:-decode or no_alloc or static}" 2>&1 | tail -5
for cfg in "7b 16 0 512" "70b 32 10 512" "70b 1 10 8192"; do
  set -- $cfg
  extra=""; [ "$3" != "0" ] && extra="--layers $3"
  timeout 600 python bench.py --steps 16 --warmup 4 --cpu-budget 0 --model $1 --batch $2 --prompt $4 $extra > gpurun_out/bench_$1_$4.json 2> gpurun_out/bench_$1_$4.err || tail -5 gpurun_out/bench_$1_$4.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_$1_$4.json"))
    print("$1 B=$2 T=$4 value",d["value"],"ms/step",d["ms_per_step"],"stage",d["stage_roofline"]["frac"],"gemm",d["roofline"]["frac"],"e2e",d["e2e"]["value"])
    print("   decode",d["kernel_ms_per_step"])
except Exception as e: print("bench $1 failed",e)
PY
done
