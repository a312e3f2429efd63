#!/usr/bin/env bash
# hand-off latency experiments on N GPUs (default 2): one micro-batch (pure latency), N micro-batches with / without step chaining
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
COMMON="bench.py --gpus $N --steps 32 --warmup 6 --cpu-budget 0 --no-parity --no-configs --no-e2e"
i=0
for extra in "--micro-batches 1" "" "--no-chain" "--no-round-graph" "--no-chain --no-round-graph" "--nccl-handoff"; do
  i=$((i+1))
  timeout 300 $TR $COMMON $extra > gpurun_out/handoff_$i.json 2> gpurun_out/handoff_$i.err || tail -5 gpurun_out/handoff_$i.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/handoff_$i.json").read().strip().splitlines()[-1])
    print("[$extra] value", d["value"], "ms/step", d["ms_per_step"], "launch", d["config"]["launch"], "handoff", d["config"]["handoff"], d["config"]["round_graph_error"])
except Exception as e: print("[$extra] failed", e)
PY
done
timeout 300 python bench.py --steps 32 --warmup 6 --cpu-budget 0 --no-e2e > gpurun_out/handoff_1gpu.json 2>/dev/null
python - <<PY
import json
d=json.load(open("gpurun_out/handoff_1gpu.json")); print("[1 gpu same box] ms/step", d["ms_per_step"])
PY
