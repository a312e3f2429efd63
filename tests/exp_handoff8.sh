#!/usr/bin/env bash
# N-GPU hand-off variants, short runs: default (stream barrier per step), --chain, one more micro-batch than stages
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
COMMON="bench.py --gpus $N --steps 32 --warmup 6 --cpu-budget 0 --no-parity --no-configs --no-e2e"
i=0
for extra in "" "--no-chain" "--micro-batches $((N+1))"; do
  i=$((i+1))
  timeout 300 $TR $COMMON $extra > gpurun_out/handoff8_$i.json 2> gpurun_out/handoff8_$i.err || tail -5 gpurun_out/handoff8_$i.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/handoff8_$i.json").read().strip().splitlines()[-1])
    print("[$extra] value", d["value"], "ms/step", d["ms_per_step"], "solo", d["stage_step_alone_ms"], "vs_slowest", d["pipeline_vs_slowest_stage"])
except Exception as e: print("[$extra] failed", e)
PY
done
