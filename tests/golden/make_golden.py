#!/usr/bin/env python
"""Generate tests/golden/*.npz by executing the UNMODIFIED reference from /root/reference.

Run from the repo root in the build container (the GPU box has no /root/reference):

    python -m tests.golden.make_golden [span|greedy|all]

The reference files are imported where they lie through oracle/ref_shim.py (two import shims, three
crash repairs — see that file); fp32, CPU.  Inputs and weights are regenerated from seeds by
``distributed_llm_inference_b200.utils.synthetic`` so only OUTPUTS are stored.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from distributed_llm_inference_b200.client import SequentialClient  # noqa: E402
from oracle import ref_shim  # noqa: E402
from tests.golden.cases import GREEDY_CASES, SPAN_CASES  # noqa: E402
from tests.golden.runner import greedy_case_weights, run_span_case  # noqa: E402


def ref_make_span(shp, weights, attn_implementation, rope):
    cfg = ref_shim.make_config(shp.hidden_size, shp.intermediate_size, shp.num_attention_heads,
                               shp.num_key_value_heads, attn_implementation, shp.rope_theta)
    blk = ref_shim.make_block(cfg, len(weights), rope=rope)
    ref_shim.load_layer_weights(blk, weights)

    def span(generation_id, hidden, past_key_value=None, cache_position=None, output_hidden_states=False,
             attention_mask=None):
        return ref_shim.block_forward(blk, generation_id, hidden, past_key_value, cache_position,
                                      output_hidden_states, attention_mask=attention_mask)

    return span


def main():
    only = sys.argv[1] if len(sys.argv) > 1 else "all"      # "span" | "greedy" | "all"
    torch.set_grad_enabled(False)
    meta = dict(torch=torch.__version__, reference="Dylan102938/distributed-llm-inference @ /root/reference",
                dtype="float32")
    names = set(sys.argv[2:])                                # optional: only these span cases
    for case in (SPAN_CASES if only in ("all", "span") else ()):
        if names and case.name not in names:
            continue
        out = run_span_case(case, ref_make_span, ref_shim.make_cache)
        path = os.path.join(HERE, f"span_{case.name}.npz")
        np.savez_compressed(path, prefill=out["prefill"], decode=out["decode"],
                            meta=np.array(repr({**meta, **case.__dict__})))
        print("wrote", path, out["prefill"].shape, out["decode"].shape)

    for case in (GREEDY_CASES if only in ("all", "greedy") else ()):
        shp, span_w, cw = greedy_case_weights(case, torch.float32)
        spans = [ref_make_span(shp, w, case.attn_implementation, case.rope) for w in span_w]
        caches = [ref_shim.make_cache(case.window, case.sinks) for _ in spans]
        client = SequentialClient(cw["embed_tokens"], cw["norm"], cw["lm_head"], spans, caches)
        from distributed_llm_inference_b200.utils import synthetic as syn

        prompt = syn.make_prompt(shp, case.batch, case.prompt)
        toks, logs = client.greedy("g0", prompt, case.n_new)
        L = torch.stack(logs, dim=1)                      # [B, n_new, V]
        top = L.topk(5, dim=-1)
        path = os.path.join(HERE, f"greedy_{case.name}.npz")
        np.savez_compressed(
            path, tokens=toks.numpy().astype(np.int64),
            logits_strided=L[:, :, :: case.vocab_stride].numpy().astype(np.float32),
            top5_values=top.values.numpy().astype(np.float32), top5_indices=top.indices.numpy().astype(np.int64),
            meta=np.array(repr({**meta, **case.__dict__})))
        margin = (top.values[..., 0] - top.values[..., 1])
        print("wrote", path, "tokens", toks[0, :8].tolist(), "min margin", margin.min().item(),
              "median", margin.median().item())


if __name__ == "__main__":
    main()
