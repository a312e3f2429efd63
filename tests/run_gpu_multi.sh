#!/usr/bin/env bash
# multi-GPU checks: N = $1 ranks on one box.  $2.. = extra bench args for the second (config-specific) run.
N=${1:-2}; shift
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 900 $TR bench.py --gpus $N --steps 16 --warmup 4 --cpu-budget 0 > gpurun_out/multi_default_$N.json 2> gpurun_out/multi_default_$N.err || tail -20 gpurun_out/multi_default_$N.err
if [ $# -gt 0 ]; then
  timeout 900 $TR bench.py --gpus $N --steps 16 --warmup 4 --cpu-budget 0 "$@" > gpurun_out/multi_cfg_$N.json 2> gpurun_out/multi_cfg_$N.err || tail -20 gpurun_out/multi_cfg_$N.err
fi
if [ -n "${CFG2:-}" ]; then
  timeout 900 $TR bench.py --gpus $N --steps 16 --warmup 4 --cpu-budget 0 $CFG2 > gpurun_out/multi_cfg2_$N.json 2> gpurun_out/multi_cfg2_$N.err || tail -20 gpurun_out/multi_cfg2_$N.err
fi
python - <<PY
import json
for f in ("gpurun_out/multi_default_$N.json", "gpurun_out/multi_cfg_$N.json", "gpurun_out/multi_cfg2_$N.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable:", e); continue
    print(f, "\n  value", d["value"], d["unit"], "ms/step", d["ms_per_step"], "stage", d["stage_roofline"]["frac"], "handoff", d["config"]["handoff"],
          "\n  workload", d["config"]["workload"][:110], "\n  prefill tok/s", d["prefill_tokens_per_sec"], d.get("prefill_pipeline"), "\n  e2e", d["e2e"])
PY
