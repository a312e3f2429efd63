"""Experiment (not a test): 70B-shaped 10-layer stage, one 8192-token prompt prefilled in chunks of C tokens on one GPU.
Per chunk size: the time of every chunk (CUDA events) and what an N-stage pipeline of such stages would take,
sum(t) + (N - 1) * mean(t) as a first-order model (every stage does the same work one chunk later)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
from distributed_llm_inference_b200.utils import synthetic as syn

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
shape = syn.SHAPES["70b"]
L = int(os.environ.get("EXP_LAYERS", "10"))
T = 8192
blk = bench.build_span(shape, range(L), dev, max_positions=T + 64)
x = syn.make_hidden(shape, 1, T, seed=5, device=dev)
sizes = [int(s) for s in os.environ.get("EXP_CHUNKS", "512,1024,2048,4096").split(",")]
for C in sizes:
    best = None
    for rep in range(3):
        cache = PartialLlamaSinkCache(1 << 20, 4, page_tokens=16, capacity_tokens=T + 256)
        evs = []
        for c0 in range(0, T, C):
            n = min(C, T - c0)
            cp = torch.arange(c0, c0 + n, device=dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            blk(f"g{C}_{rep}", x[:, c0:c0 + n].contiguous(), past_key_value=cache, cache_position=cp)
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        ts = [a.elapsed_time(b) for a, b in evs]
        if best is None or sum(ts) < sum(best):
            best = ts
        del cache
    tot, mean = sum(best), sum(best) / len(best)
    print(f"chunk {C:5d}: {len(best):2d} chunks, stage total {tot:7.2f} ms, first {best[0]:6.2f} last {best[-1]:6.2f}; "
          f"pipeline model 8 stages {tot + 7 * mean:7.1f} ms, 4 stages (x2 layers) {2 * (tot + 3 * mean):7.1f} ms", flush=True)
