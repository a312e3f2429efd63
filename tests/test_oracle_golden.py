"""Pin the CPU oracle: (1) against the committed fixtures produced by the reference itself
(tests/golden/make_golden.py); (2) where /root/reference is present, against the reference run live."""
import os

import numpy as np
import pytest
import torch

from distributed_llm_inference_b200.utils import synthetic as syn
from oracle import ref_shim
from oracle import span_oracle as so
from tests.golden.cases import GREEDY_CASES, SPAN_CASES
from tests.golden.runner import greedy_case_weights, run_span_case

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def oracle_make_span(shp, weights, attn_implementation, rope):
    shape = so.SpanShape(shp.hidden_size, shp.intermediate_size, shp.num_attention_heads,
                         shp.num_key_value_heads, shp.head_dim, shp.rope_theta)
    layers = [so.LayerWeights(*(w[k].float() for k in syn.LINEARS + syn.NORMS)) for w in weights]
    return so.OracleSpan(shape, layers, attn_implementation, rope)


@pytest.mark.parametrize("case", SPAN_CASES, ids=lambda c: c.name)
def test_oracle_matches_reference_fixture(case):
    gold = np.load(os.path.join(GOLD, f"span_{case.name}.npz"))
    out = run_span_case(case, oracle_make_span, so.SinkCacheOracle)
    # fp32 vs fp32, same library: differences come only from BLAS blocking on another host
    np.testing.assert_allclose(out["prefill"], gold["prefill"], rtol=0, atol=2e-5)
    np.testing.assert_allclose(out["decode"], gold["decode"], rtol=0, atol=2e-5)


@pytest.mark.parametrize("case", GREEDY_CASES, ids=lambda c: c.name)
def test_oracle_greedy_matches_reference_fixture(case):
    gold = np.load(os.path.join(GOLD, f"greedy_{case.name}.npz"))
    shp, span_w, cw = greedy_case_weights(case, torch.float32)
    spans = [oracle_make_span(shp, w, case.attn_implementation, case.rope) for w in span_w]
    caches = [so.SinkCacheOracle(case.window, case.sinks) for _ in spans]
    client = so.OracleClient(cw["embed_tokens"], cw["norm"], cw["lm_head"], spans, caches)
    prompt = syn.make_prompt(shp, case.batch, case.prompt)
    # teacher-force the reference's tokens so that a 1e-6 tie-break cannot desynchronise the streams
    forced = torch.from_numpy(gold["tokens"])
    toks, logs = client.greedy("g0", prompt, case.n_new, forced=forced)
    L = torch.stack(logs, dim=1)
    np.testing.assert_allclose(L[:, :, :: case.vocab_stride].numpy(), gold["logits_strided"], rtol=0, atol=5e-5)
    margin = gold["top5_values"][..., 0] - gold["top5_values"][..., 1]
    agree = toks.numpy() == gold["tokens"]
    assert agree[margin > 1e-4].all()


@pytest.mark.skipif(not ref_shim.available(), reason="/root/reference not present on this machine")
@pytest.mark.parametrize("impl,rope", [("eager", "float"), ("sdpa", "literal")])
def test_oracle_matches_reference_live(impl, rope):
    """Two concurrent generations with different batch sizes + eviction, reference run live."""
    shp = syn.SHAPES["test_gqa"]
    W = syn.make_span_weights(shp, [5, 6], dtype=torch.float32, norm_jitter=0.2)
    cfg = ref_shim.make_config(shp.hidden_size, shp.intermediate_size, shp.num_attention_heads,
                               shp.num_key_value_heads, impl)
    blk = ref_shim.make_block(cfg, 2, rope=rope)
    ref_shim.load_layer_weights(blk, W)
    osp = oracle_make_span(shp, W, impl, rope)
    rc, oc = ref_shim.make_cache(10, 2), so.SinkCacheOracle(10, 2)
    lens = {"a": 0, "b": 0}
    for step, (gid, B, L) in enumerate([("a", 1, 5), ("b", 3, 7), ("a", 1, 1), ("b", 3, 1), ("a", 1, 1),
                                        ("b", 3, 2), ("a", 1, 1), ("b", 3, 1), ("a", 1, 1), ("a", 1, 1),
                                        ("a", 1, 1), ("a", 1, 1)]):
        x = syn.make_hidden(shp, B, L, seed=50 + step, dtype=torch.float32)
        cp = torch.arange(lens[gid], lens[gid] + L)
        r = ref_shim.block_forward(blk, gid, x, rc, cache_position=cp, output_hidden_states=True)
        o = osp.forward(gid, x, past_key_value=oc, cache_position=cp, output_hidden_states=True)
        lens[gid] += L
        assert len(r) == len(o) == 2 and len(r[1]) == len(o[1]) == 3
        assert torch.equal(r[0], o[0]), (step, (r[0] - o[0]).abs().max())
        assert rc.get_seq_length(0, gid) == oc.get_seq_length(0, gid)
        assert rc._seen_tokens[gid] == oc._seen_tokens[gid]


def test_oracle_error_behaviour():
    """cache.py:24-30,51-52,71-72 raise ValueError when the generation id is missing."""
    c = so.SinkCacheOracle(8, 2)
    with pytest.raises(ValueError):
        c.get_seq_length(0, None)
    with pytest.raises(ValueError):
        c.update(torch.zeros(1, 1, 1, 4), torch.zeros(1, 1, 1, 4), 0, cache_kwargs={})
    assert c.get_seq_length(0, "unknown") == 0
