"""Thin test-side helpers that call the C ABI (through distributed_llm_inference_b200._ffi) on torch tensors."""
from __future__ import annotations

import ctypes as C

import torch

from distributed_llm_inference_b200 import _ffi

lib = _ffi.lib


def stream():
    return _ffi.current_stream()


def rmsnorm(x, w, eps, pre_scale=1.0, want_resid=False):
    rows, h = x.shape
    y = torch.empty(rows, h, dtype=torch.bfloat16, device=x.device)
    resid = torch.empty(rows, h, dtype=torch.float32, device=x.device) if want_resid else None
    _ffi.check(lib.dli_rmsnorm(x.data_ptr(), int(x.dtype == torch.float32), w.data_ptr(), y.data_ptr(),
                               resid.data_ptr() if resid is not None else None, rows, h, eps, pre_scale, stream()))
    return (y, resid) if want_resid else y


def gemm(x, w, epilogue, out=None):
    """x [m, k] bf16, w [n or 2n, k] bf16."""
    m, k = x.shape
    n = w.shape[0] // 2 if epilogue == _ffi.EPI_SILU_MUL_BF16 else w.shape[0]
    if out is None:
        dt = torch.float32 if epilogue in (_ffi.EPI_STORE_F32, _ffi.EPI_ADD_F32) else torch.bfloat16
        out = torch.zeros(m, n, dtype=dt, device=x.device)
    _ffi.check(lib.dli_gemm(x.data_ptr(), x.stride(0), w.data_ptr(), out.data_ptr(), out.stride(0), m, n, k,
                            epilogue, stream()))
    return out


class PagedKV:
    """A page pool + one generation's block table, built by hand for operator tests."""

    def __init__(self, n_layers, batch, kv_heads, head_dim, max_tokens, page_tokens=16, device="cuda", seed=0):
        self.pt, self.B, self.KV, self.hd = page_tokens, batch, kv_heads, head_dim
        self.max_pages = -(-max_tokens // page_tokens)
        n_pages = batch * self.max_pages + 3
        g = torch.Generator().manual_seed(seed)
        perm = torch.randperm(n_pages, generator=g)[: batch * self.max_pages]      # scattered physical pages
        self.block_table = perm.view(batch, self.max_pages).to(torch.int32).to(device)
        self.k = torch.zeros(n_layers, n_pages, kv_heads, page_tokens, head_dim, dtype=torch.bfloat16, device=device)
        self.v = torch.zeros_like(self.k)
        self.seq_lens = torch.zeros(batch, dtype=torch.int32, device=device)
        self.n_pages = n_pages

    def desc(self, max_ctx):
        return _ffi.KvDesc(self.k.data_ptr(), self.v.data_ptr(), self.k.stride(0) * 2, self.pt, self.n_pages,
                           self.block_table.data_ptr(), self.seq_lens.data_ptr(), self.max_pages, max_ctx)

    def gather(self, layer, total):
        """-> K, V as [B, KV, total, hd] (what the reference's cache tensors would hold)."""
        pos = torch.arange(total, device=self.k.device)
        pg = self.block_table.long()[:, pos // self.pt]
        sl = (pos % self.pt)[None, :].expand(self.B, total)
        return self.k[layer][pg, :, sl].transpose(1, 2), self.v[layer][pg, :, sl].transpose(1, 2)

    def fill(self, layer, k, v):
        """k, v: [B, KV, T, hd] -> pages; sets seq_lens = T."""
        T = k.shape[2]
        pos = torch.arange(T, device=self.k.device)
        pg = self.block_table.long()[:, pos // self.pt]
        sl = (pos % self.pt)[None, :].expand(self.B, T)
        self.k[layer][pg, :, sl] = k.transpose(1, 2).to(torch.bfloat16)
        self.v[layer][pg, :, sl] = v.transpose(1, 2).to(torch.bfloat16)
        self.seq_lens.fill_(T)


def rope_kv_append(qkv, cos, sin, kv: PagedKV, layer, batch, q_len, H, KV, hd, zero_after=False):
    q = torch.empty(batch * q_len, H * hd, dtype=torch.bfloat16, device=qkv.device)
    d = kv.desc(int(kv.seq_lens.max().item()) + q_len)
    _ffi.check(lib.dli_rope_kv_append(qkv.data_ptr(), int(zero_after), cos.data_ptr(), sin.data_ptr(), q.data_ptr(),
                                      C.byref(d), layer, batch, q_len, H, KV, hd, stream()))
    return q


def decode_attention(q, kv: PagedKV, layer, batch, H, KV, hd, max_ctx):
    out = torch.empty(batch, H * hd, dtype=torch.bfloat16, device=q.device)
    nbytes = lib.dli_decode_attention_scratch_bytes(batch, H, hd, max_ctx)
    scratch = torch.empty(nbytes // 4, dtype=torch.float32, device=q.device)
    d = kv.desc(max_ctx)
    _ffi.check(lib.dli_decode_attention(q.data_ptr(), out.data_ptr(), C.byref(d), layer, batch, H, KV, hd,
                                        scratch.data_ptr(), nbytes, stream()))
    return out


def prefill_attention(q, kv: PagedKV, layer, batch, q_len, H, KV, hd, causal, max_ctx):
    out = torch.empty(batch * q_len, H * hd, dtype=torch.bfloat16, device=q.device)
    d = kv.desc(max_ctx)
    _ffi.check(lib.dli_prefill_attention(q.data_ptr(), out.data_ptr(), C.byref(d), layer, batch, q_len, H, KV, hd,
                                         int(causal), stream()))
    return out


def err_report(name, got, ref):
    got, ref = got.float().cpu(), ref.float().cpu()
    d = (got - ref).abs()
    scale = ref.abs().max().item() + 1e-12
    idx = torch.nonzero(d == d.max())[0].tolist() if d.numel() else []
    bad = (d > 0.05 * scale)
    msg = (f"{name}: shape {tuple(ref.shape)} max_abs_err {d.max().item():.4e} mean_abs_err {d.mean().item():.4e} "
           f"ref_max {scale:.4e} at {idx} bad_frac {bad.float().mean().item():.4f}")
    if bad.any() and got.dim() == 2:
        rows = torch.nonzero(bad.any(1)).flatten()
        cols = torch.nonzero(bad.any(0)).flatten()
        msg += f" bad_rows[{rows.numel()}] {rows[:8].tolist()}.. bad_cols[{cols.numel()}] {cols[:8].tolist()}.."
    return msg, d.max().item(), scale
