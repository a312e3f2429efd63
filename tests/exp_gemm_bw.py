"""Experiment (not a test): effective weight-streaming bandwidth of dli_gemm for different weight shapes."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from distributed_llm_inference_b200 import _ffi
from tests import gpu_ops as ops

dev = "cuda"
M = int(os.environ.get("EXP_M", "16"))
shapes = [(12288, 4096), (4096, 4096), (22016, 4096), (4096, 11008)]
if os.environ.get("EXP_70B"):
    shapes = [(10240, 8192), (8192, 8192), (57344, 8192), (8192, 28672)]
for (n, k) in shapes:
    nbuf = max(2, int(600e6 // (n * k * 2)) + 1)            # cycle through > 126 MB (L2) of distinct weights
    ws = [torch.randn(n, k, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
    x = torch.randn(M, k, device=dev, dtype=torch.bfloat16)
    out = torch.zeros(M, n, device=dev, dtype=torch.float32)
    for w in ws:
        ops.gemm(x, w, _ffi.EPI_ADD_F32, out=out)
    torch.cuda.synchronize()
    reps = 6
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for w in ws:
            ops.gemm(x, w, _ffi.EPI_ADD_F32, out=out)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * nbuf)
    print(f"m={M} n={n:6d} k={k:6d} {n*k*2/1e6:7.1f} MB  {us:7.2f} us  {n*k*2/us/1e3:7.1f} GB/s  MIN_KB={os.environ.get('DLI_MIN_KB','4')} FILL_X2={os.environ.get('DLI_FILL_X2','4')}")
    del ws
