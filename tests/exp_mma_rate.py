"""Experiment (not a test): time per tcgen05.mma in a single-CTA accumulation chain, by MMA-N."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from distributed_llm_inference_b200 import _ffi
from tests import gpu_ops as ops
K = 32768
for m in (16, 32, 64, 128):
    for n in (128, 128 * 148):
        x = torch.randn(m, K, device="cuda").to(torch.bfloat16)
        w = torch.randn(n, K, device="cuda").to(torch.bfloat16)
        out = torch.zeros(m, n, device="cuda", dtype=torch.bfloat16)
        for _ in range(3):
            ops.gemm(x, w, _ffi.EPI_STORE_BF16, out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            ops.gemm(x, w, _ffi.EPI_STORE_BF16, out=out)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 100
        mmas = K // 16
        print(f"MMA N={m:3d} tiles={n//128:4d}: {us:8.1f} us per launch, {us * 1e3 / mmas:7.1f} ns per MMA (~{us * 1.9e3 / mmas:6.0f} clk @1.9GHz), "
              f"weights {n*K*2/1e6:.0f} MB -> {n*K*2/us/1e3:.0f} GB/s")
