"""Experiment (not a test): decode step time of the 7B span under different launch modes."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from distributed_llm_inference_b200.utils import synthetic as syn
from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache

B = int(os.environ.get("EXP_B", "16")); CTX = 512; STEPS = 40
dev = torch.device("cuda", 0)
shape = syn.SHAPES["7b"]
blk = bench.build_span(shape, list(range(shape.num_hidden_layers)), dev)
cache = PartialLlamaSinkCache(1 << 20, 4, capacity_tokens=B * 1024)


def fill(gid):
    x = syn.make_hidden(shape, B, CTX, seed=1, device=dev)
    blk(gid, x, past_key_value=cache)
    cache.reserve(gid, B, 1000)


def loop(gid, stream=None, steps=STEPS):
    x = syn.make_hidden(shape, B, 1, seed=2, device=dev)
    pos = cache.get_seq_length(0, gid)
    ctxm = torch.cuda.stream(stream) if stream is not None else torch.cuda.stream(torch.cuda.current_stream())
    with ctxm:
        for _ in range(5):
            x = blk(gid, x, past_key_value=cache, cache_position=torch.tensor([pos], device=dev))[0]; pos += 1
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            x = blk(gid, x, past_key_value=cache, cache_position=torch.tensor([pos], device=dev))[0]; pos += 1
        e1.record()
        t_cpu = time.perf_counter() - t0
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, t_cpu / steps * 1e3


fill("a")
print(f"B={B} PDL={'off' if os.environ.get('DLI_NO_PDL') == '1' else 'on'}")
ms, cpu = loop("a")
print(f"default stream : {ms:.3f} ms/step (cpu enqueue {cpu:.3f} ms/step)")
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
ms, cpu = loop("a", s)
print(f"side stream    : {ms:.3f} ms/step (cpu enqueue {cpu:.3f} ms/step)")

# CUDA graph of one decode step (static input / position tensors; seq_lens advanced on the device inside the graph)
gen = cache.generation("a")
xs = syn.make_hidden(shape, B, 1, seed=2, device=dev)
cp = torch.tensor([gen.length], device=dev)
g = torch.cuda.CUDAGraph()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(2):
        blk("a", xs, past_key_value=cache, cache_position=cp); cp += 1
torch.cuda.current_stream().wait_stream(s)
torch.cuda.synchronize()
with torch.cuda.graph(g):
    ys = blk("a", xs, past_key_value=cache, cache_position=cp)[0]
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(3):
    g.replay(); cp += 1
torch.cuda.synchronize()
e0.record()
for _ in range(STEPS):
    g.replay(); cp += 1
e1.record()
torch.cuda.synchronize()
print(f"cuda graph     : {e0.elapsed_time(e1) / STEPS:.3f} ms/step")
