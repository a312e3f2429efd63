"""Drive one golden case through ANY implementation of the span interface.

An implementation is a pair of factories:
  make_span(case_or_shape, layer_weights: list[dict], attn_implementation, rope) -> callable
      span(generation_id, hidden, past_key_value=cache, cache_position=cp, output_hidden_states=False)
      returning the reference's tuple (model.py:76)
  make_cache(window, sinks) -> cache object accepted by that span
The same driver produces the fixtures from the reference (make_golden.py), checks the oracle
(test_oracle_golden.py) and checks the CUDA path (test_span_gpu.py).
"""
from __future__ import annotations

from typing import Callable, Dict, List

import numpy as np
import torch

from distributed_llm_inference_b200.utils import synthetic as syn

from .cases import GreedyCase, SpanCase


def span_case_inputs(case: SpanCase, dtype, device="cpu"):
    shp = syn.SHAPES[case.shape]
    x0 = syn.make_hidden(shp, case.batch, case.prefill, seed=7, dtype=dtype, device=device)
    steps = [syn.make_hidden(shp, case.batch, 1, seed=100 + t, dtype=dtype, device=device)
             for t in range(case.prefill, case.prefill + case.n_decode)]
    return shp, x0, steps


def span_case_weights(case: SpanCase, dtype, device="cpu"):
    shp = syn.SHAPES[case.shape]
    ids = list(range(case.first_layer, case.first_layer + case.n_layers))
    # always drawn by the CPU generator (what the fixtures were made with), then moved
    ws = syn.make_span_weights(shp, ids, dtype=dtype, device="cpu", norm_jitter=case.norm_jitter)
    return [{k: v.to(device) for k, v in w.items()} for w in ws]


def case_mask(case: SpanCase, length: int, device="cpu"):
    """2-D attention mask ``[B, length]`` of the case (1 = token, 0 = padding inside the prompt), or None."""
    if not case.pad:
        return None
    m = torch.ones(case.batch, length, dtype=torch.long)
    for b, n in enumerate(case.pad):
        if n:
            if case.pad_side == "left":
                m[b, :n] = 0
            else:
                m[b, case.prefill - n:case.prefill] = 0
    return m.to(device)


def run_span_case(case: SpanCase, make_span: Callable, make_cache: Callable, dtype=torch.float32,
                  device="cpu") -> Dict[str, np.ndarray]:
    shp, x0, steps = span_case_inputs(case, dtype, device)
    weights = span_case_weights(case, dtype, device)
    span = make_span(shp, weights, case.attn_implementation, case.rope)
    cache = make_cache(case.window, case.sinks)
    out: Dict[str, np.ndarray] = {}
    mk = {} if not case.pad else None          # keyword only when the case has a mask (older factories take none)
    with torch.no_grad():
        kw = {} if mk is not None else {"attention_mask": case_mask(case, case.prefill, device)}
        y = span("g0", x0, past_key_value=cache, cache_position=None, **kw)[0]
        out["prefill"] = y.float().cpu().numpy()
        dec: List[np.ndarray] = []
        if case.chunks:
            pos = case.prefill
            for i, n in enumerate(case.chunks):
                xt = syn.make_hidden(shp, case.batch, n, seed=1000 + i, dtype=dtype, device=device)
                cp = torch.arange(pos, pos + n, device=x0.device)
                kw = {} if mk is not None else {"attention_mask": case_mask(case, pos + n, device)}
                y = span("g0", xt, past_key_value=cache, cache_position=cp, **kw)[0]
                dec.append(y.float().cpu().numpy())
                pos += n
            out["decode"] = np.concatenate(dec, axis=1)          # [B, sum(chunks), h]
            return out
        for i, xt in enumerate(steps):
            cp = torch.tensor([case.prefill + i], device=x0.device)
            kw = {} if mk is not None else {"attention_mask": case_mask(case, case.prefill + i + 1, device)}
            y = span("g0", xt, past_key_value=cache, cache_position=cp, **kw)[0]
            dec.append(y.float().cpu().numpy())
        out["decode"] = np.stack(dec, axis=0) if dec else np.zeros((0,), np.float32)
    return out


def greedy_case_weights(case: GreedyCase, dtype, device="cpu"):
    shp = syn.SHAPES[case.shape]
    spans = [[{k: v.to(device) for k, v in w.items()} for w in
              syn.make_span_weights(shp, list(range(a, b)), dtype=dtype, device="cpu")] for a, b in case.spans]
    make = syn.make_decidable_client_weights if case.client == "decidable" else syn.make_client_weights
    client = {k: v.to(device) for k, v in make(shp, dtype=dtype, device="cpu").items()}
    return shp, spans, client
