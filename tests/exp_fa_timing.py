"""Experiment (not a test): per-phase cycle counters of one CTA of the tcgen05 attention kernel."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
dbg = torch.zeros(16, dtype=torch.int64, device="cuda")
os.environ["DLI_FA_DBG_PTR"] = hex(dbg.data_ptr())
from tests import gpu_ops as ops
B, H, KV, hd, L = 1, 64, 8, 128, 8192
kv = ops.PagedKV(1, B, KV, hd, L + 64, page_tokens=int(os.environ.get("EXP_PT", "16")))
K = torch.randn(B, KV, L, hd, device="cuda").to(torch.bfloat16); V = torch.randn_like(K)
kv.fill(0, K, V); kv.seq_lens.fill_(0)
q = torch.randn(B * L, H * hd, device="cuda").to(torch.bfloat16)
for _ in range(2):
    out = ops.prefill_attention(q, kv, 0, B, L, H, KV, hd, True, max_ctx=L)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); out = ops.prefill_attention(q, kv, 0, B, L, H, KV, hd, True, max_ctx=L); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
d = dbg.tolist(); n = d[5]
print(f"kernel {ms:.3f} ms ({4*L*L*H*hd/2/ms/1e9:.0f} TFLOP/s causal); CTA with {n} key tiles, cycles per tile:")
if os.environ.get("DLI_FA_V1"):
    names = ["wait s_full", "S load", "compute", "wait pv_done (+rescale)", "P store + fences + arrive"]
    mnames = ["QK warp: wait k_full", "QK warp: wait s_empty", "QK warp: issue+rest", "PV warp: wait v_full", "PV warp: wait p_full", "PV warp: issue+rest"]
else:
    names = ["wait s_full", "S load", "mask + max", "row-max exchange barrier", "rescale + exp + P store + arrive"]
    mnames = ["MMA warp: wait v_full", "MMA warp: wait p_full A", "MMA warp: wait k_full", "MMA warp: wait p_full B", "MMA warp: issue + rest"]
for name, v in zip(names, d[:5]):
    print(f"  {name:34s} {v / max(n,1):8.0f}")
print(f"  total per tile {sum(d[:5]) / max(n,1):8.0f}")
for name, v in zip(mnames, d[8:14]):
    print(f"  {name:34s} {v / max(n,1):8.0f}")
