#!/usr/bin/env bash
# A/B of the per-round CUDA graph on N GPUs (same box): bench with and without --no-round-graph
N=${1:-2}; shift
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29566"
for rep in ${REPS:-1 2}; do
for flag in "" "--no-round-graph"; do
  timeout 600 $TR bench.py --gpus $N --steps 32 --warmup 4 --cpu-budget 0 $flag "$@" > gpurun_out/rg.json 2> gpurun_out/rg.err || tail -20 gpurun_out/rg.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/rg.json").read().strip().splitlines()[-1])
print("flag '$flag':", d["value"], "tok/s", d["ms_per_step"], "ms/round", d["config"]["launch"], d["config"].get("round_graph_error"), "e2e", d["e2e"]["value"])
PY
done
done
