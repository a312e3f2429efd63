"""Experiment: per-phase cycle counters of the int8-weight GEMM's converter warps (W8_TIMING build)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
dbg = torch.zeros(16, dtype=torch.int64, device="cuda")
os.environ["DLI_W8_DBG_PTR"] = hex(dbg.data_ptr())
from distributed_llm_inference_b200 import _ffi
from tests import gpu_ops as ops
M, n, k = 16, 12288, 4096
x = torch.randn(M, k, device="cuda", dtype=torch.bfloat16)
out = torch.zeros(M, n, device="cuda")
qs = [torch.randint(0, 256, (n, k), device="cuda", dtype=torch.uint8) for _ in range(6)]
sc = torch.rand(n, device="cuda")
for rep in range(3):
    for q in qs:
        _ffi.check(_ffi.lib.dli_gemm_w8(x.data_ptr(), k, q.data_ptr(), sc.data_ptr(), out.data_ptr(), n, M, n, k, _ffi.EPI_ADD_F32, ops.stream()))
torch.cuda.synchronize()
d = dbg.tolist()
for name, o in (("warp 2 lane 0", 0), ("warp 9 lane 0", 8)):
    kb = max(1, d[o + 4])
    print(f"{name}: k-blocks {kb}; cycles per k-block: wait raw {d[o]/kb:.0f}, wait slot {d[o+1]/kb:.0f}, load+widen+store {d[o+2]/kb:.0f}, fence+arrive {d[o+3]/kb:.0f}")
