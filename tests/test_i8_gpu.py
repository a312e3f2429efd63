"""int8 activations on int8 weights (LLM.int8, the arithmetic of the reference's optimised mode, utils/model.py:93-123):
the activation quantiser bit for bit against the oracle's restatement, the tcgen05 kind::i8 decode GEMM against the
oracle's arithmetic (outliers included), and a quantised span against the fp32 oracle span whose linear layers apply the
same activation quantisation.  (oracle/quant_oracle.py is restated from bitsandbytes' published algorithm: unpinned.)"""
import ctypes as C

import pytest
import torch

from distributed_llm_inference_b200 import _ffi
from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
from distributed_llm_inference_b200.utils import synthetic as syn
from oracle import quant_oracle as qo
from oracle import span_oracle as so
from tests import gpu_ops as ops
from tests.test_oracle_golden import oracle_make_span
from tests.test_span_gpu import _check, gpu_make_span
from tests.test_w8_gpu import _quantize_gpu

pytestmark = pytest.mark.gpu
DEV = "cuda"


class ActI8:
    """A dli_act_i8 carved out of one torch buffer, with torch views of its arrays."""

    def __init__(self, m, k, planes):
        n = _ffi.lib.dli_act_i8_bytes(m, k, planes)
        self.buf = torch.zeros(n, dtype=torch.uint8, device=DEV)
        self.d = _ffi.ActI8()
        _ffi.check(_ffi.lib.dli_act_i8_carve(self.buf.data_ptr(), n, m, k, planes, C.byref(self.d)))
        self.m, self.k, self.planes, self.kb = m, k, planes, (k + 127) // 128

    def _view(self, ptr, nbytes, dtype):
        off = ptr - self.buf.data_ptr()
        return self.buf[off:off + nbytes].view(dtype)

    def plane(self, i):
        m, k, kb = self.m, self.k, self.kb
        qp, sp, kp = ((self.d.q, self.d.scale, self.d.ksum), (self.d.q2, self.d.scale2, self.d.ksum2))[i]
        return (self._view(qp, m * k, torch.int8).view(m, k).cpu(), self._view(sp, m * 4, torch.float32).cpu(),
                self._view(kp, m * kb * 4, torch.int32).view(m, kb).cpu())


def _make_x(m, k, g, n_outliers):
    x = torch.randn(m, k, generator=g)
    x = x.clamp_(-4.5, 4.5)
    for i in range(n_outliers):                                   # a few entries at or above the threshold, some sharing a column
        r = int(torch.randint(0, m, (1,), generator=g))
        c = 17 if i % 3 == 0 else int(torch.randint(0, k, (1,), generator=g))
        x[r, c] = (6.0 + 1.37 * i) * (1 if i % 2 else -1)
    return x.to(torch.bfloat16)


def _quantize_act(x, threshold, planes):
    m, k = x.shape
    a = ActI8(m, k, planes)
    xd = x.to(DEV)
    _ffi.check(_ffi.lib.dli_quantize_act_i8(xd.data_ptr(), k, m, k, threshold, C.byref(a.d), ops.stream()))
    torch.cuda.synchronize()
    return a


def _block_sums(codes, k):
    pad = (-k) % 128
    return torch.nn.functional.pad(codes.to(torch.int32), (0, pad)).view(codes.shape[0], -1, 128).sum(dim=2).to(torch.int32)


@pytest.mark.parametrize("m", [1, 16, 33, 128])
@pytest.mark.parametrize("k", [64, 1408, 4096, 11008])
def test_quantize_act_bit_exact(m, k):
    g = torch.Generator().manual_seed(m * 131 + k)
    x = _make_x(m, k, g, n_outliers=min(12, m * 3))
    x[0] = 0                                                       # an all-zero row: scales 0, codes 0
    a = _quantize_act(x, 5.0, 2)
    r1, rs1, r2, rs2 = qo.quantize_act_rowwise(x, 5.0)
    for (q, sc, ks), (rq, rs) in zip((a.plane(0), a.plane(1)), ((r1, rs1), (r2, rs2))):
        assert torch.equal(q, rq)
        assert torch.equal(sc, rs)
        assert torch.equal(ks, _block_sums(rq, k))
    assert int((r2 != 0).sum()) > 0 or m == 1
    # no threshold, one plane: plain row-wise absmax quantisation
    b = _quantize_act(x, 0.0, 1)
    q0, s0, k0 = b.plane(0)
    rq0, rs0, _, _ = qo.quantize_act_rowwise(x, 0.0)
    assert torch.equal(q0, rq0) and torch.equal(s0, rs0) and torch.equal(k0, _block_sums(rq0, k))


@pytest.mark.parametrize("m", [1, 16, 33, 128])
@pytest.mark.parametrize("n,k", [(256, 512), (12288, 4096), (4096, 11008), (1000, 1408)])
@pytest.mark.parametrize("epi", [_ffi.EPI_STORE_F32, _ffi.EPI_ADD_F32, _ffi.EPI_STORE_BF16, _ffi.EPI_SILU_MUL_BF16])
def test_gemm_i8_matches_oracle(m, n, k, epi):
    g = torch.Generator().manual_seed(m * 7 + n + k + epi)
    rows = 2 * n if epi == _ffi.EPI_SILU_MUL_BF16 else n
    w = (0.02 * torch.randn(rows, k, generator=g)).to(torch.bfloat16)
    x = _make_x(m, k, g, n_outliers=min(10, 2 * m))
    planes, thr = (1, 0.0) if m == 33 else (2, 5.0)                # this batch size: one plane, no threshold
    q, s_w = _quantize_gpu(w.to(DEV))
    rq, rs = qo.quantize_rowwise(w)
    a = _quantize_act(x, thr, planes)
    ref = qo.linear_llm_int8_rowwise(x, rq, rs, thr)
    f32 = epi in (_ffi.EPI_STORE_F32, _ffi.EPI_ADD_F32)
    base = torch.randn(m, n, generator=g) if epi == _ffi.EPI_ADD_F32 else torch.zeros(m, n)
    out = (base.clone() if f32 else torch.zeros(m, n, dtype=torch.bfloat16)).to(DEV)
    if epi == _ffi.EPI_STORE_F32:
        out.fill_(7.0)                                             # must be overwritten, not accumulated into
    _ffi.check(_ffi.lib.dli_gemm_i8(C.byref(a.d), q.data_ptr(), s_w.data_ptr(), out.data_ptr(), n, m, n, k, epi, ops.stream()))
    if epi == _ffi.EPI_SILU_MUL_BF16:
        ref = torch.nn.functional.silu(ref[:, :n]) * ref[:, n:]
    elif epi == _ffi.EPI_ADD_F32:
        ref = ref + base
    msg, err, scale = ops.err_report(f"gemm_i8 m{m} n{n} k{k} epi{epi}", out, ref)
    tol = 1e-3 if f32 else 6e-3                                    # exact integer products; fp32 scaling; bf16 output rounding
    assert err <= tol * scale + 1e-5, msg
    # against the unquantised product: the mode's own error (reported, loosely bounded)
    exact = x.float() @ qo.dequantize_rowwise(rq, rs).t()
    if epi == _ffi.EPI_STORE_F32 and planes == 2:
        e = (out.cpu() - exact).abs().max() / exact.abs().max()
        assert e < 0.05, f"int8 activations: {e:.3e} of the output scale"


def _act_quant_linear(threshold):
    def linear(x, w, bias=None):
        rows = x.numel() // x.shape[-1]
        if rows > 128:                                             # chunks of more than 128 tokens run on 16-bit activations
            return torch.nn.functional.linear(x, w, bias)
        x2 = x.reshape(rows, x.shape[-1]).float()
        c1, s1, c2, s2 = qo.quantize_act_rowwise(x2, threshold)
        xq = c1.float() * s1[:, None] + c2.float() * s2[:, None]
        return (xq @ w.float().t()).reshape(*x.shape[:-1], w.shape[0])
    return linear


@pytest.mark.parametrize("shape_name,B,L1,L2,threshold", [("test_hd128", 2, 40, 3, 5.0), ("test_gqa", 2, 200, 30, 3.0)])
def test_int8_activation_span_matches_oracle(shape_name, B, L1, L2, threshold):
    from tests.test_w8_gpu import _dequantised_weights

    shp = syn.SHAPES[shape_name]
    W = syn.make_span_weights(shp, [0, 1, 2], norm_jitter=0.1)
    from distributed_llm_inference_b200.utils.model import convert_to_optimized_block

    blk = convert_to_optimized_block(gpu_make_span(shp, W, "eager", "float"), quantize=True, threshold=threshold)   # the reference's call
    osp = oracle_make_span(shp, _dequantised_weights(W), "eager", "float")
    gc, oc = PartialLlamaSinkCache(4096, 4, capacity_tokens=4096), so.SinkCacheOracle(4096, 4)
    so.LINEAR = _act_quant_linear(threshold)
    try:
        pos = 0
        for step, L in enumerate([L1, L2, 1, 1]):
            x = syn.make_hidden(shp, B, L, seed=40 + step)
            cp = torch.arange(pos, pos + L)
            ref = osp.forward("g", x.float(), past_key_value=oc, cache_position=cp)[0]
            got = blk("g", x.to(DEV), past_key_value=gc, cache_position=cp.to(DEV))[0]
            _check(f"i8 span chunk {step} ({L} tokens)", got.float().cpu().numpy(), ref.numpy())
            pos += L
        graph = blk.graphed_decode("g", gc, max_ctx=pos + 8)
        for step in range(3):
            xt = syn.make_hidden(shp, B, 1, seed=60 + step)
            ref = osp.forward("g", xt.float(), past_key_value=oc, cache_position=torch.tensor([pos + step]))[0]
            _check(f"i8 graphed step {step}", graph.step(xt.to(DEV)).float().cpu().numpy(), ref.numpy())
    finally:
        so.LINEAR = torch.nn.functional.linear
    # back to 16-bit activations on the same span object
    blk.quantize_w8(activations="bf16")
    x = syn.make_hidden(shp, B, 1, seed=99)
    blk("g", x.to(DEV), past_key_value=gc, cache_position=torch.tensor([pos + 3], device=DEV))
    assert blk._a8_applied is None
