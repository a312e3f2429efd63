#!/usr/bin/env bash
# Round-end check as the driver runs it: GPU test suite, smoke, default bench, reference arm; records under gpurun_out/.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu --no-header -p no:cacheprovider > gpurun_out/r02t_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -n 3 gpurun_out/r02t_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02t_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/r02t_smoke.log
timeout 900 python bench.py > gpurun_out/r02t_bench_1gpu.json 2> gpurun_out/r02t_bench_1gpu.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02t_bench_reference.json 2> gpurun_out/r02t_bench_reference.err; echo "ref exit $?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02t_bench_1gpu.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_step","gpu_launches","clocks")}, d["roofline"], d["e2e"], d.get("decode_int8_weights",{}).get("value"))
r=json.loads(open("gpurun_out/r02t_bench_reference.json").read().strip().splitlines()[-1])
print({k:r.get(k) for k in ("impl","value","ms_per_step","cpu_baseline")})
PY
