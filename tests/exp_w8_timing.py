"""Experiment: per-phase cycle counters of the int8-weight GEMM's converter warps (W8_TIMING build)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
dbg = torch.zeros(16, dtype=torch.int64, device="cuda")
os.environ["DLI_W8_DBG_PTR"] = hex(dbg.data_ptr())
from distributed_llm_inference_b200 import _ffi
from tests import gpu_ops as ops
M, n, k = 16, int(os.environ.get("EXP_N", "12288")), int(os.environ.get("EXP_K", "4096"))
silu = n < 0                      # negative n: gate|up with the SiLU epilogue
n = abs(n)
rows = 2 * n if silu else n
epi = _ffi.EPI_SILU_MUL_BF16 if silu else _ffi.EPI_ADD_F32
x = torch.randn(M, k, device="cuda", dtype=torch.bfloat16)
out = torch.zeros(M, n, device="cuda", dtype=torch.bfloat16 if silu else torch.float32)
qs = [torch.randint(0, 256, (rows, k), device="cuda", dtype=torch.uint8) for _ in range(6)]
sc = torch.rand(rows, device="cuda")
for rep in range(3):
    for q in qs:
        _ffi.check(_ffi.lib.dli_gemm_w8(x.data_ptr(), k, q.data_ptr(), sc.data_ptr(), out.data_ptr(), n, M, n, k, epi, ops.stream()))
torch.cuda.synchronize()
d = dbg.tolist()
for name, o in (("first converter warp", 0), ("last converter warp", 8)):
    kb = max(1, d[o + 4])
    print(f"{name}: k-blocks {kb}; cycles per k-block: wait raw {d[o]/kb:.0f}, wait slot {d[o+1]/kb:.0f}, load+widen+store {d[o+2]/kb:.0f}, fence+arrive {d[o+3]/kb:.0f}")
