"""Weight-only int8 (convert_to_optimized_block(quantize=True); the weight format of the reference's Linear8bitLt,
utils/model.py:93-123): quantisation kernel bit for bit against the oracle's restatement, the decode GEMM against the
oracle's arithmetic, and a quantised span -- short and long prefill, eager and graphed decode -- against the fp32 oracle
span run on the dequantised weights.  (The oracle for this mode is restated from bitsandbytes' published algorithm and
is unpinned: oracle/quant_oracle.py.)"""

import pytest
import torch

from distributed_llm_inference_b200 import _ffi
from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
from distributed_llm_inference_b200.utils import synthetic as syn
from distributed_llm_inference_b200.utils.model import convert_to_optimized_block
from oracle import quant_oracle as qo
from oracle import span_oracle as so
from tests import gpu_ops as ops
from tests.test_oracle_golden import oracle_make_span
from tests.test_span_gpu import _check, gpu_make_span

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _quantize_gpu(w):
    n, k = w.shape
    q = torch.empty(n, k, dtype=torch.uint8, device=DEV)
    s = torch.empty(n, dtype=torch.float32, device=DEV)
    _ffi.check(_ffi.lib.dli_quantize_rows_w8(w.data_ptr(), n, k, q.data_ptr(), s.data_ptr(), ops.stream()))
    return q, s


@pytest.mark.parametrize("n,k", [(7, 64), (300, 1408), (4096, 4096)])
def test_quantize_rows_bit_exact_and_roundtrip(n, k):
    g = torch.Generator().manual_seed(n + k)
    w = (0.02 * torch.randn(n, k, generator=g)).to(torch.bfloat16)
    w[0] = 0                                                       # an all-zero row: scale 0, codes 128
    q, s = _quantize_gpu(w.to(DEV))
    rq, rs = qo.quantize_rowwise(w)
    assert torch.equal(q.cpu().to(torch.int16) - 128, rq.to(torch.int16))
    assert torch.equal(s.cpu(), rs)
    back = torch.empty(n, k, dtype=torch.bfloat16, device=DEV)
    _ffi.check(_ffi.lib.dli_dequantize_rows_w8(q.data_ptr(), s.data_ptr(), n, k, back.data_ptr(), ops.stream()))
    assert torch.equal(back.cpu(), qo.dequantize_rowwise(rq, rs).to(torch.bfloat16))


@pytest.mark.parametrize("m", [1, 16, 33, 128])
@pytest.mark.parametrize("n,k", [(256, 512), (12288, 4096), (4096, 11008), (1000, 1408)])
@pytest.mark.parametrize("epi", [_ffi.EPI_STORE_F32, _ffi.EPI_ADD_F32, _ffi.EPI_STORE_BF16, _ffi.EPI_SILU_MUL_BF16])
def test_gemm_w8_matches_oracle(m, n, k, epi):
    g = torch.Generator().manual_seed(m * 7 + n + k + epi)
    rows = 2 * n if epi == _ffi.EPI_SILU_MUL_BF16 else n
    w = (0.02 * torch.randn(rows, k, generator=g)).to(torch.bfloat16)
    x = torch.randn(m, k, generator=g).to(torch.bfloat16)
    q, s = _quantize_gpu(w.to(DEV))
    rq, rs = qo.quantize_rowwise(w)
    ref = qo.linear_w8(x, rq, rs)
    f32 = epi in (_ffi.EPI_STORE_F32, _ffi.EPI_ADD_F32)
    base = torch.randn(m, n, generator=g) if epi == _ffi.EPI_ADD_F32 else torch.zeros(m, n)
    out = (base.clone() if f32 else torch.zeros(m, n, dtype=torch.bfloat16)).to(DEV)
    if epi == _ffi.EPI_STORE_F32:
        out.fill_(7.0)                                             # must be overwritten, not accumulated into
    _ffi.check(_ffi.lib.dli_gemm_w8(x.to(DEV).data_ptr(), k, q.data_ptr(), s.data_ptr(), out.data_ptr(), n, m, n, k, epi,
                                    ops.stream()))
    if epi == _ffi.EPI_SILU_MUL_BF16:
        ref = torch.nn.functional.silu(ref[:, :n]) * ref[:, n:]
    elif epi == _ffi.EPI_ADD_F32:
        ref = ref + base
    msg, err, scale = ops.err_report(f"gemm_w8 m{m} n{n} k{k} epi{epi}", out, ref)
    tol = 1e-3 if f32 else 6e-3                                    # exact products, fp32 accumulate; bf16 output rounding
    assert err <= tol * scale + 1e-5, msg


def _dequantised_weights(W):
    """The weights the GPU path multiplies by: every fused matrix quantised row-wise and widened again (oracle side)."""
    Wd = []
    for w in W:                                                    # the oracle sees the weights the GPU path multiplies by
        d = dict(w)
        qkv = torch.cat([w["self_attn.q_proj"], w["self_attn.k_proj"], w["self_attn.v_proj"]])
        for name, t in (("qkv", qkv), ("self_attn.o_proj", w["self_attn.o_proj"]),
                        ("gu", torch.cat([w["mlp.gate_proj"], w["mlp.up_proj"]])), ("mlp.down_proj", w["mlp.down_proj"])):
            dq = qo.dequantize_rowwise(*qo.quantize_rowwise(t))
            if name == "qkv":
                nq, nkv = w["self_attn.q_proj"].shape[0], w["self_attn.k_proj"].shape[0]
                d["self_attn.q_proj"], d["self_attn.k_proj"], d["self_attn.v_proj"] = dq[:nq], dq[nq:nq + nkv], dq[nq + nkv:]
            elif name == "gu":
                ni = w["mlp.gate_proj"].shape[0]
                d["mlp.gate_proj"], d["mlp.up_proj"] = dq[:ni], dq[ni:]
            else:
                d[name] = dq
        Wd.append(d)
    return Wd


@pytest.mark.parametrize("shape_name,B,L1,L2", [("test_hd128", 2, 40, 3), ("test_gqa", 2, 200, 150)])
def test_quantised_span_matches_oracle_on_dequantised_weights(shape_name, B, L1, L2):
    shp = syn.SHAPES[shape_name]
    W = syn.make_span_weights(shp, [0, 1, 2], norm_jitter=0.1)
    blk = convert_to_optimized_block(gpu_make_span(shp, W, "eager", "float"), quantize=True, activations="bf16")
    assert all(layer.w8 is not None and layer.self_attn.q_proj.weight.numel() == 0 for layer in blk.layers)
    osp = oracle_make_span(shp, _dequantised_weights(W), "eager", "float")
    gc, oc = PartialLlamaSinkCache(4096, 4, capacity_tokens=4096), so.SinkCacheOracle(4096, 4)
    pos = 0
    for step, L in enumerate([L1, L2, 1, 1]):
        x = syn.make_hidden(shp, B, L, seed=40 + step)
        cp = torch.arange(pos, pos + L)
        ref = osp.forward("g", x.float(), past_key_value=oc, cache_position=cp)[0]
        got = blk("g", x.to(DEV), past_key_value=gc, cache_position=cp.to(DEV))[0]
        _check(f"w8 span chunk {step} ({L} tokens)", got.float().cpu().numpy(), ref.numpy())
        pos += L
    graph = blk.graphed_decode("g", gc, max_ctx=pos + 8)
    for step in range(3):
        xt = syn.make_hidden(shp, B, 1, seed=60 + step)
        ref = osp.forward("g", xt.float(), past_key_value=oc, cache_position=torch.tensor([pos + step]))[0]
        _check(f"w8 graphed step {step}", graph.step(xt.to(DEV)).float().cpu().numpy(), ref.numpy())
