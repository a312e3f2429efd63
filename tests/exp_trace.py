"""Experiment (not a test): in-kernel timeline of one graph-replayed decode step.

    DLI_TRACE=1 bash distributed_llm_inference_b200/csrc/build.sh
    DLI_LIB=$PWD/distributed_llm_inference_b200/csrc/libdli_b200_trace.so python tests/exp_trace.py

Every CTA stamps %globaltimer at entry, when griddepcontrol.wait returns, and at exit (common.cuh, DLI_TRACE).
The script groups the stamps into launches and prints, per launch: when the first CTA became resident, when the
dependency wait returned, when the last CTA left, and the gap to the next launch's post-wait start.
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import bench
from distributed_llm_inference_b200 import _ffi
from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
from distributed_llm_inference_b200.utils import synthetic as syn

B = int(os.environ.get("EXP_B", "16"))
CTX = int(os.environ.get("EXP_CTX", "512"))
MODEL = os.environ.get("EXP_MODEL", "7b")
LAYERS = int(os.environ.get("EXP_LAYERS", "0"))
dev = torch.device("cuda", 0)
shape = syn.SHAPES[MODEL]
n_layers = LAYERS or shape.num_hidden_layers
blk = bench.build_span(shape, list(range(n_layers)), dev)
cache = PartialLlamaSinkCache(1 << 20, 4, capacity_tokens=B * (CTX + 512))
x = syn.make_hidden(shape, B, CTX, seed=1, device=dev)
blk("a", x, past_key_value=cache)
g = blk.graphed_decode("a", cache, CTX + 64)
for _ in range(5):
    g.step()
torch.cuda.synchronize()

CAP = 1 << 20
buf = torch.zeros(2 + 2 * CAP, dtype=torch.int64, device=dev)
buf[1] = CAP
lib = _ffi.lib
lib.dli_debug_trace.argtypes = [C.c_void_p]
lib.dli_debug_trace.restype = C.c_int
assert lib.dli_debug_trace(buf.data_ptr()) == 0
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
g.step()
e1.record()
torch.cuda.synchronize()
print(f"traced step: {e0.elapsed_time(e1):.3f} ms")
n = int(buf[0].item())
ev = buf[2:2 + 2 * n].cpu().numpy().astype(np.uint64).reshape(n, 2)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.save(os.path.join(ROOT, "gpurun_out", f"trace_{MODEL}_B{B}.npy"), ev)
tag, t = ev[:, 0], ev[:, 1].astype(np.int64)
kid = (tag >> np.uint64(56)).astype(int)
ph = ((tag >> np.uint64(52)) & np.uint64(15)).astype(int)
t0 = t.min()
t = (t - t0) / 1e3  # us
NAMES = {1: "gemm", 2: "rmsnorm", 3: "rope", 4: "attn", 5: "combine", 6: "cast", 7: "attn_gqa"}
# group into launches: walk events sorted by the time of phase 1 (dependency resolved) -- launches are serialised by
# the dependency chain, so phase-1 stamps of consecutive launches do not interleave
order = np.argsort(t, kind="stable")
launches = []
cur = None
p1 = [(t[i], kid[i]) for i in order if ph[i] == 1]
# cluster phase-1 stamps: a new launch starts when the kernel id changes or there is a > 1.2 us hole
groups = []
for tt, k in p1:
    if groups and groups[-1]["kid"] == k and tt - groups[-1]["w1"] < 1.2:
        groups[-1]["w1"] = tt
        groups[-1]["n"] += 1
    else:
        groups.append({"kid": k, "w0": tt, "w1": tt, "n": 1})
# attach entry / exit stamps: entry = earliest phase-0 of that kernel id after the previous same-id group's entry window
for gi, gr in enumerate(groups):
    nxt = groups[gi + 1]["w0"] if gi + 1 < len(groups) else 1e18
    sel2 = (kid == gr["kid"]) & (ph == 2) & (t >= gr["w0"]) & (t <= nxt + 50)
    # exits belonging to this launch: CTAs of this kernel that exit after its wait and before the same kernel id waits again
    later_same = [g2["w0"] for g2 in groups[gi + 1:] if g2["kid"] == gr["kid"]]
    lim = later_same[0] if later_same else 1e18
    sel2 &= t < lim
    gr["end"] = t[sel2].max() if sel2.any() else float("nan")
    gr["end_first"] = t[sel2].min() if sel2.any() else float("nan")
    prev_same = [g2["w0"] for g2 in groups[:gi] if g2["kid"] == gr["kid"]]
    lo = prev_same[-1] if prev_same else -1.0
    sel0 = (kid == gr["kid"]) & (ph == 0) & (t > lo - 1e-9) & (t <= gr["w1"])
    # entries of this launch are those after the previous same-id launch's wait-return
    gr["entry"] = t[sel0].min() if sel0.any() else float("nan")
print(f"{n} events, {len(groups)} launches")
print("  #  kernel        entry    wait0    wait1   end_first   end_last   dur(wait0->end)  idle_to_next_wait0")
L = int(os.environ.get("EXP_SHOW", "40"))
for gi, gr in enumerate(groups[:L]):
    nxt = groups[gi + 1]["w0"] if gi + 1 < len(groups) else float("nan")
    name = NAMES.get(gr["kid"] & 15, "?") + (f"/e{gr['kid'] >> 4}" if (gr["kid"] & 15) == 1 else "")
    print(f"{gi:3d}  {name:10s} {gr['entry']:8.2f} {gr['w0']:8.2f} {gr['w1']:8.2f} {gr['end_first']:10.2f} {gr['end']:10.2f} "
          f"{gr['end'] - gr['w0']:12.2f} {nxt - gr['end']:14.2f}   ctas {gr['n']}")
# per-class summary over the whole step
import collections
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for gi, gr in enumerate(groups[:-1]):
    name = NAMES.get(gr["kid"] & 15, "?") + (f"/e{gr['kid'] >> 4}" if (gr["kid"] & 15) == 1 else "")
    a = agg[name]
    a[0] += 1
    a[1] += gr["end"] - gr["w0"]
    a[2] += groups[gi + 1]["w0"] - gr["end"]
print("class        launches   busy_us(wait0->end)   gap_us(end->next wait0)")
for k, (c, b_, g_) in agg.items():
    print(f"{k:12s} {c:5d} {b_:14.1f} {g_:14.1f}   per launch {b_ / c:7.2f} {g_ / c:7.2f}")
