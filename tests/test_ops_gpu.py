"""Operator-level parity: each sm_100a kernel, called through the C ABI, against the oracle's restatement of the
reference lines it replaces (same seeded inputs, fp32 CPU/GPU torch math).  Tolerances are stated per test."""
import math

import pytest
import torch

from distributed_llm_inference_b200 import _ffi
from oracle import span_oracle as so
from tests import gpu_ops as ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _randn(*shape, seed, scale=1.0, dtype=torch.bfloat16):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dtype).to(DEV)


# ------------------------------------------------------------------------------------------------ RMSNorm
@pytest.mark.parametrize("rows,h", [(1, 512), (5, 768), (33, 4096), (16, 8192), (300, 5120)])
@pytest.mark.parametrize("f32_in", [False, True])
@pytest.mark.parametrize("pre_scale", [1.0, 2.0])
def test_rmsnorm(rows, h, f32_in, pre_scale):
    x = _randn(rows, h, seed=1, scale=1.7, dtype=torch.float32 if f32_in else torch.bfloat16)
    w = (1.0 + 0.2 * _randn(h, seed=2, dtype=torch.float32)).to(torch.bfloat16)
    eps = 1e-6
    y, resid = ops.rmsnorm(x, w, eps, pre_scale, want_resid=True)
    ref = so.rmsnorm(x.float().cpu() * pre_scale, w.float().cpu(), eps)          # modules.py:162 / :179 (s = 2)
    # one bf16 rounding of the output: 2^-9 relative, plus fp32 reduction-order noise
    torch.testing.assert_close(y.float().cpu(), ref, rtol=6e-3, atol=1e-4)
    assert torch.equal(resid.cpu(), x.float().cpu())


# ------------------------------------------------------------------------------------------------ GEMM
GEMM_SHAPES = [
    # (m, n, k)  decode-like (swap-AB, split-K) ...
    (1, 256, 512), (4, 768, 768), (16, 2304, 768), (17, 1024, 4096), (32, 4096, 4096), (64, 1536, 1024),
    (100, 512, 1408), (128, 384, 512), (16, 4096, 11008), (8, 200, 512),
    # ... and prefill-like (token-major tiles)
    (129, 384, 512), (200, 1024, 768), (512, 2304, 768), (1000, 4096, 1024), (300, 200, 512),
    # ... enough tiles that every CTA pair of the persistent kernel loops (74 pairs), ragged last token tile
    (1537, 2560, 1024), (4096, 5120, 512),
]


@pytest.mark.parametrize("m,n,k", GEMM_SHAPES)
@pytest.mark.parametrize("epi", [_ffi.EPI_STORE_F32, _ffi.EPI_ADD_F32, _ffi.EPI_STORE_BF16])
def test_gemm(m, n, k, epi):
    x = _randn(m, k, seed=3)
    w = _randn(n, k, seed=4, scale=0.05)
    ref = x.float() @ w.float().t()                                              # nn.Linear, modules.py:61-63,112,181
    if epi == _ffi.EPI_ADD_F32:
        base = _randn(m, n, seed=5, dtype=torch.float32)
        out = ops.gemm(x, w, epi, out=base.clone())
        ref = ref + base                                                          # residual add, modules.py:173,182
    else:
        out = ops.gemm(x, w, epi)
    torch.cuda.synchronize()
    msg, err, scale = ops.err_report(f"gemm m{m} n{n} k{k} epi{epi}", out, ref)
    # exact bf16 products, fp32 accumulation (order differs): 1e-3 of the output scale; bf16 store adds 2^-9
    tol = (1e-3 if epi != _ffi.EPI_STORE_BF16 else 6e-3) * scale
    assert err <= tol, msg


@pytest.mark.parametrize("m,n,k", [(129, 128, 512), (300, 1408, 512), (512, 2048, 768), (1000, 704, 1024),
                                   (1, 1408, 512), (16, 2048, 768), (100, 704, 1024), (8, 200, 512), (128, 11008, 4096),
                                   (2304, 2816, 512)])
def test_gemm_silu_mul(m, n, k):
    x = _randn(m, k, seed=6)
    w = _randn(2 * n, k, seed=7, scale=0.05)                                      # gate rows then up rows
    out = ops.gemm(x, w, _ffi.EPI_SILU_MUL_BF16)
    g, u = x.float() @ w[:n].float().t(), x.float() @ w[n:].float().t()
    ref = torch.nn.functional.silu(g) * u                                         # LlamaMLP.forward, modules.py:181
    msg, err, scale = ops.err_report(f"gemm_silu m{m} n{n} k{k}", out, ref)
    assert err <= 8e-3 * scale, msg


def test_gemm_rejects_bad_arguments():
    x = _randn(4, 512, seed=1)
    w = _randn(64, 512, seed=2)
    out = torch.zeros(4, 64, dtype=torch.float32, device=DEV)
    rc = ops.lib.dli_gemm(x.data_ptr(), 512, w.data_ptr(), out.data_ptr(), 64, 4, 64, 510, 0, ops.stream())
    assert rc == -1 and b"multiples of 8" in ops.lib.dli_last_error()
    rc = ops.lib.dli_gemm(None, 512, w.data_ptr(), out.data_ptr(), 64, 4, 64, 512, 0, ops.stream())
    assert rc == -1


# ------------------------------------------------------------------------------------------------ SiLU*up, cast
def test_silu_mul_and_cast():
    m, n = 7, 1408
    gu = _randn(m, 2 * n, seed=8, dtype=torch.float32)
    keep = gu.clone()
    act = torch.empty(m, n, dtype=torch.bfloat16, device=DEV)
    _ffi.check(ops.lib.dli_silu_mul(gu.data_ptr(), act.data_ptr(), m, n, 1, ops.stream()))
    ref = torch.nn.functional.silu(keep[:, :n]) * keep[:, n:]
    torch.testing.assert_close(act.float(), ref, rtol=6e-3, atol=1e-5)
    assert not gu.any()                                                            # accumulation buffer re-zeroed
    out = torch.empty(m, 2 * n, dtype=torch.bfloat16, device=DEV)
    _ffi.check(ops.lib.dli_cast_f32_bf16(keep.data_ptr(), out.data_ptr(), keep.numel(), ops.stream()))
    assert torch.equal(out, keep.to(torch.bfloat16))


# ------------------------------------------------------------------------------------------------ RoPE + append
@pytest.mark.parametrize("H,KV,hd", [(8, 2, 64), (4, 4, 128), (8, 1, 128)])
@pytest.mark.parametrize("B,past,L", [(2, 0, 9), (3, 21, 1), (1, 30, 37)])
@pytest.mark.parametrize("mode", ["float", "literal"])
def test_rope_kv_append(H, KV, hd, B, past, L, mode):
    kv = ops.PagedKV(2, B, KV, hd, past + L + 5)
    kv.seq_lens.fill_(past)
    n_all = (H + 2 * KV) * hd
    qkv = _randn(B * L, n_all, seed=9, dtype=torch.float32)
    pos = torch.arange(past, past + L)[None]
    cos, sin = so.rope_tables(pos, hd, 10000.0, mode, torch.float32)              # model.py:55
    cos, sin = cos[0].float().contiguous().to(DEV), sin[0].float().contiguous().to(DEV)
    work = qkv.clone()
    q = ops.rope_kv_append(work, cos, sin, kv, 1, B, L, H, KV, hd, zero_after=True)
    assert not work.any()
    x = qkv.view(B, L, n_all).cpu()
    qr = x[..., : H * hd].view(B, L, H, hd).transpose(1, 2)
    kr = x[..., H * hd: (H + KV) * hd].view(B, L, KV, hd).transpose(1, 2)
    vr = x[..., (H + KV) * hd:].view(B, L, KV, hd).transpose(1, 2)
    qe, ke = so.apply_rotary_pos_emb(qr, kr, cos.cpu()[None], sin.cpu()[None])    # modules.py:17-20
    torch.testing.assert_close(q.view(B, L, H, hd).transpose(1, 2).float().cpu(), qe, rtol=6e-3, atol=1e-5)
    K, V = kv.gather(1, past + L)
    torch.testing.assert_close(K[:, :, past:].float().cpu(), ke, rtol=6e-3, atol=1e-5)
    torch.testing.assert_close(V[:, :, past:].float().cpu(), vr, rtol=6e-3, atol=1e-5)
    assert not kv.k[0].any() and not K[:, :, :past].any()                         # nothing else touched


# ------------------------------------------------------------------------------------------------ attention
def _attn_ref(q, K, V, causal, past):
    """modules.py:87-105 in fp32 on bf16-rounded inputs.  q [B, H, L, hd]; K, V [B, KV, T, hd]."""
    B, H, L, hd = q.shape
    K, V = so.repeat_kv(K, H // K.shape[1]), so.repeat_kv(V, H // V.shape[1])
    s = torch.matmul(q, K.transpose(2, 3)) / math.sqrt(hd)
    if causal:
        T = K.shape[2]
        j = torch.arange(T)[None, :]
        i = past + torch.arange(L)[:, None]
        s = s.masked_fill(j > i, float("-inf"))
    p = torch.softmax(s, dim=-1, dtype=torch.float32)
    return torch.matmul(p, V).transpose(1, 2).reshape(B, L, H * hd)


@pytest.mark.parametrize("H,KV,hd", [(8, 2, 64), (4, 4, 128), (8, 1, 128), (12, 12, 64), (4, 2, 128)])
@pytest.mark.parametrize("B,T", [(1, 1), (2, 17), (3, 200), (1, 1000), (16, 513)])
def test_decode_attention(H, KV, hd, B, T):
    kv = ops.PagedKV(1, B, KV, hd, T + 3, seed=T)
    K, V = _randn(B, KV, T, hd, seed=10), _randn(B, KV, T, hd, seed=11)
    kv.fill(0, K, V)
    kv.seq_lens.fill_(T - 1)                                                       # T_b = seq_lens + 1
    q = _randn(B, H * hd, seed=12)
    out = ops.decode_attention(q, kv, 0, B, H, KV, hd, max_ctx=T)
    ref = _attn_ref(q.view(B, H, 1, hd).float().cpu(), K.float().cpu(), V.float().cpu(), False, T - 1)[:, 0]
    msg, err, scale = ops.err_report(f"decode_attn H{H} KV{KV} hd{hd} B{B} T{T}", out, ref)
    assert err <= 8e-3 * scale + 1e-3, msg                                         # bf16 output rounding


@pytest.mark.parametrize("H,KV,hd", [(4, 4, 128), (8, 1, 128), (8, 4, 128), (16, 4, 64), (64, 8, 128)])
@pytest.mark.parametrize("pt", [8, 16, 32, 64, 24])
@pytest.mark.parametrize("B,T", [(3, 333), (1, 4500)])
def test_decode_attention_ragged_pages(H, KV, hd, pt, B, T):
    """Ragged contexts over several page sizes (24 does not tile a stage -> the generic kernel), enough context
    for many splits, and NaN in every slot no sequence owns: nothing outside [0, T_b) may reach the output."""
    kv = ops.PagedKV(1, B, KV, hd, T + 3, page_tokens=pt, seed=T + pt)
    K, V = _randn(B, KV, T, hd, seed=20), _randn(B, KV, T, hd, seed=21)
    kv.k.fill_(float("nan"))
    kv.v.fill_(float("nan"))
    kv.fill(0, K, V)
    lens = [max(1, T - 97 * b - (b % 2)) for b in range(B)]                        # T_b per sequence
    pos = torch.arange(T, device="cuda")
    pg = kv.block_table.long()[:, pos // pt]
    sl = (pos % pt)[None, :].expand(B, T)
    dead = pos[None, :] >= torch.tensor(lens, device="cuda")[:, None]              # slots past each sequence's end
    kk, vv = kv.k[0][pg, :, sl], kv.v[0][pg, :, sl]                                # [B, T, KV, hd]
    kk[dead] = float("nan")
    vv[dead] = float("nan")
    kv.k[0][pg, :, sl] = kk
    kv.v[0][pg, :, sl] = vv
    kv.seq_lens.copy_(torch.tensor([n - 1 for n in lens], dtype=torch.int32))
    q = _randn(B, H * hd, seed=22)
    out = ops.decode_attention(q, kv, 0, B, H, KV, hd, max_ctx=T)
    assert torch.isfinite(out.float()).all()
    for b in range(B):
        n = lens[b]
        ref = _attn_ref(q[b:b + 1].view(1, H, 1, hd).float().cpu(), K[b:b + 1, :, :n].float().cpu(),
                        V[b:b + 1, :, :n].float().cpu(), False, n - 1)[:, 0]
        msg, err, scale = ops.err_report(f"decode_attn ragged H{H} KV{KV} hd{hd} pt{pt} b{b} T{n}", out[b:b + 1], ref)
        assert err <= 8e-3 * scale + 1e-3, msg


@pytest.mark.parametrize("H,KV,hd", [(8, 2, 64), (4, 4, 128), (8, 1, 128), (4, 2, 128)])
@pytest.mark.parametrize("B,past,L", [(1, 0, 1), (2, 0, 9), (1, 0, 64), (2, 0, 200), (1, 37, 100), (3, 5, 65), (1, 0, 1000),
                                      (2, 0, 256), (1, 100, 300), (2, 130, 515), (1, 700, 257),
                                      # >= 2048 queries: the two-tile TS-mode kernel (attn_prefill_tc2.cu)
                                      (1, 0, 2048), (1, 333, 2300), (2, 0, 2100)])
@pytest.mark.parametrize("causal", [True, False])
def test_prefill_attention(H, KV, hd, B, past, L, causal):
    T = past + L
    kv = ops.PagedKV(1, B, KV, hd, T + 3, seed=L)
    K, V = _randn(B, KV, T, hd, seed=13), _randn(B, KV, T, hd, seed=14)
    kv.fill(0, K, V)
    kv.seq_lens.fill_(past)
    q = _randn(B * L, H * hd, seed=15)
    out = ops.prefill_attention(q, kv, 0, B, L, H, KV, hd, causal, max_ctx=T)
    qr = q.view(B, L, H, hd).transpose(1, 2).float().cpu()
    ref = _attn_ref(qr, K.float().cpu(), V.float().cpu(), causal, past).reshape(B * L, H * hd)
    msg, err, scale = ops.err_report(f"prefill_attn H{H} KV{KV} hd{hd} B{B} past{past} L{L} causal{causal}", out, ref)
    # P is rounded to bf16 as the tensor-core operand (as modules.py:96 does) and the output to bf16
    assert err <= 1.5e-2 * scale + 1e-3, msg
