"""Experiment (not a test): weight-streaming rate of the int8-weight decode GEMM (dli_gemm_w8) next to the bf16 one."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from distributed_llm_inference_b200 import _ffi
from tests import gpu_ops as ops

dev = "cuda"
M = int(os.environ.get("EXP_M", "16"))
shapes = [(12288, 4096), (4096, 4096), (4096, 11008), (-11008, 4096)]    # negative n: gate|up with the SiLU epilogue
only_w8 = bool(os.environ.get("EXP_ONLY_W8"))
for (n, k) in shapes:
    silu = n < 0
    n = abs(n)
    rows = 2 * n if silu else n
    epi = _ffi.EPI_SILU_MUL_BF16 if silu else _ffi.EPI_ADD_F32
    nbuf = max(2, int(400e6 // (rows * k)) + 1)
    x = torch.randn(M, k, device=dev, dtype=torch.bfloat16)
    out = torch.zeros(M, n, device=dev, dtype=torch.bfloat16 if silu else torch.float32)
    qs = [torch.randint(0, 256, (rows, k), device=dev, dtype=torch.uint8) for _ in range(nbuf)]
    sc = torch.rand(rows, device=dev) * 1e-3

    def run_w8():
        for q in qs:
            _ffi.check(_ffi.lib.dli_gemm_w8(x.data_ptr(), k, q.data_ptr(), sc.data_ptr(), out.data_ptr(), n, M, n, k,
                                            epi, ops.stream()))

    def timeit(fn, reps=6):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (reps * nbuf)

    us8 = timeit(run_w8)
    # int8 activations (two planes, threshold 5) quantised once; the kind::i8 GEMM alone
    import ctypes as C
    nb = _ffi.lib.dli_act_i8_bytes(M, k, 2)
    abuf = torch.zeros(nb, dtype=torch.uint8, device=dev)
    ad = _ffi.ActI8()
    _ffi.check(_ffi.lib.dli_act_i8_carve(abuf.data_ptr(), nb, M, k, 2, C.byref(ad)))
    _ffi.check(_ffi.lib.dli_quantize_act_i8(x.data_ptr(), k, M, k, 5.0, C.byref(ad), ops.stream()))
    ad1 = _ffi.ActI8(ad.q, ad.scale, ad.ksum, None, None, None)

    def run_i8(a):
        for q in qs:
            _ffi.check(_ffi.lib.dli_gemm_i8(C.byref(a), q.data_ptr(), sc.data_ptr(), out.data_ptr(), n, M, n, k, epi, ops.stream()))

    usi2 = timeit(lambda: run_i8(ad))
    usi1 = timeit(lambda: run_i8(ad1))
    line = f"m={M} n={n:6d} k={k:6d} {'silu' if silu else 'add '} int8 {rows*k/1e6:6.1f} MB {us8:7.2f} us {rows*k/us8/1e3:7.1f} GB/s | i8 2 planes {usi2:7.2f} us {rows*k/usi2/1e3:7.1f} GB/s, 1 plane {usi1:7.2f} us"
    if not only_w8:
        ws = [torch.randn(rows, k, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
        us16 = timeit(lambda: [ops.gemm(x, w, epi, out=out) for w in ws])
        line += f" | bf16 {us16:7.2f} us {rows*k*2/us16/1e3:7.1f} GB/s | speed-up {us16/us8:.2f}"
        del ws
    print(line, flush=True)
    del qs
