"""Parity at BASELINE.json's full layer shapes (Llama-2-7B and -70B widths), where the CPU oracle is only affordable
on small token counts.  Two kinds of checks:

* against the oracle: 2 full-width layers on a short chunk + decode steps (a few hundred tokens of fp32 CPU work);
* size-independent properties at the bench's own batch and prompt sizes, where the oracle would take minutes:
    - a prompt prefilled in one chunk and in two chunks gives the same output and the same cache
      (the two runs take different GEMM tile counts and the second chunk attends to a cached past);
    - prefill(T) + decode(1) equals the last row of prefill(T + 1) (decode kernels against prefill kernels);
    - a replayed decode graph equals the eager step.
  "Same" is bf16-level: the two sides round at the same places but reduce in different orders.
"""
import numpy as np
import pytest
import torch

from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
from distributed_llm_inference_b200.utils import synthetic as syn
from oracle import span_oracle as so
from tests.test_oracle_golden import oracle_make_span
from tests.test_span_gpu import _check, gpu_make_span

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _close(name, a, b, rel=2e-2):
    a, b = a.float().cpu().numpy(), b.float().cpu().numpy()
    scale = np.abs(b).max()
    d = np.abs(a - b)
    assert np.isfinite(a).all() and np.isfinite(b).all(), f"{name}: non-finite"
    assert d.max() <= rel * scale, f"{name}: max_abs_diff {d.max():.4e} (mean {d.mean():.4e}) vs scale {scale:.4e}"
    assert d.mean() <= 0.1 * rel * scale, f"{name}: mean_abs_diff {d.mean():.4e} vs scale {scale:.4e}"


@pytest.mark.parametrize("shape_name,B,L", [("7b", 2, 160), ("13b", 2, 130), ("70b", 1, 136)])
def test_full_width_layers_match_oracle(shape_name, B, L):
    """Two layers of the real width: a chunk long enough for the CTA-pair GEMM (>= 256 tokens in total) and the
    tcgen05 attention, then decode steps through the swap-AB GEMMs and the MHA / GQA decode attention."""
    shp = syn.SHAPES[shape_name]
    W = syn.make_span_weights(shp, [0, 1], norm_jitter=0.1)
    blk = gpu_make_span(shp, W, "eager", "float")
    osp = oracle_make_span(shp, W, "eager", "float")
    gc, oc = PartialLlamaSinkCache(4096, 4, capacity_tokens=2048), so.SinkCacheOracle(4096, 4)
    x = syn.make_hidden(shp, B, L, seed=31)
    ref = osp.forward("g", x.float(), past_key_value=oc)[0]
    got = blk("g", x.to(DEV), past_key_value=gc)[0]
    _check(f"{shape_name} prefill", got.float().cpu().numpy(), ref.numpy())
    for step in range(3):
        xt = syn.make_hidden(shp, B, 1, seed=40 + step)
        cp = torch.tensor([L + step])
        ref = osp.forward("g", xt.float(), past_key_value=oc, cache_position=cp)[0]
        got = blk("g", xt.to(DEV), past_key_value=gc, cache_position=cp.to(DEV))[0]
        _check(f"{shape_name} decode {step}", got.float().cpu().numpy(), ref.numpy())


# (70b, 32, 512) is the one where the unsplit GQA decode kernel also rotates q/k and appends (rope_inline.cuh)
@pytest.mark.parametrize("shape_name,B,T", [("7b", 16, 512), ("13b", 8, 512), ("70b", 4, 512), ("70b", 1, 2048),
                                            ("70b", 32, 512)])
def test_bench_size_properties(shape_name, B, T):
    shp = syn.SHAPES[shape_name]
    W = syn.make_span_weights(shp, [0, 1], norm_jitter=0.1)
    blk = gpu_make_span(shp, W, "eager", "float")
    cache = PartialLlamaSinkCache(1 << 20, 4, capacity_tokens=4 * B * (T + 64))
    x = syn.make_hidden(shp, B, T + 1, seed=5).to(DEV)
    xt = x[:, T:T + 1].contiguous()

    # (1) one chunk vs two chunks
    one = blk("one", x[:, :T].contiguous(), past_key_value=cache)[0]
    h = T // 2 + 64                                      # uneven split: different tile counts on the two sides
    two_a = blk("two", x[:, :h].contiguous(), past_key_value=cache)[0]
    two_b = blk("two", x[:, h:T].contiguous(), past_key_value=cache, cache_position=torch.arange(h, T, device=DEV))[0]
    _close("chunked prefill, first chunk", two_a, one[:, :h])
    _close("chunked prefill, second chunk", two_b, one[:, h:])
    pos = torch.arange(T, device=DEV)

    def cached(layer, gid):                              # -> K, V [B, T, KV, hd] as the pages hold them
        g = cache.generation(gid)
        pg = g.block_table.long()[:, pos // cache.page_tokens]
        sl = (pos % cache.page_tokens)[None, :].expand(B, T)
        return cache._k[layer][pg, :, sl], cache._v[layer][pg, :, sl]

    for layer in range(2):
        (k1, v1), (k2, v2) = cached(layer, "one"), cached(layer, "two")
        _close(f"cached K layer {layer}", k2, k1, rel=1e-2)
        _close(f"cached V layer {layer}", v2, v1, rel=1e-2)

    # (2) prefill(T) + decode(1) vs prefill(T + 1)
    full = blk("full", x, past_key_value=cache)[0]
    dec = blk("one", xt, past_key_value=cache, cache_position=torch.tensor([T], device=DEV))[0]
    _close("decode step vs last prefill row", dec[:, 0], full[:, T])

    # (3) graph replay vs eager, on a generation in the same state
    graph = blk.graphed_decode("two", cache, max_ctx=T + 16)
    rep = graph.step(xt).clone()
    _close("graph replay vs eager step", rep[:, 0], dec[:, 0], rel=1e-2)   # the two caches differ by bf16 ulps (1)
    assert cache.get_seq_length(0, "one") == cache.get_seq_length(0, "two") == T + 1
