"""Run the UNMODIFIED reference files from /root/reference — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Used only to (a) validate ``oracle/span_oracle.py`` and (b) generate the committed fixtures in
``tests/golden/`` (``tests/golden/make_golden.py``).  /root/reference does not exist on the GPU box, so
nothing that runs there may import this module; ``available()`` tells callers whether it can work.

No reference source is copied: the reference package is imported from where it lies.  Two import
shims are needed because the reference pins transformers==4.43.3 and this image has 5.5.0
(SURVEY section 8c):

  shim 1  ``transformers.cache_utils.SinkCache`` was removed in 5.x; a stub with the two helper
          methods the reference inherits (``_rotate_half``, ``_apply_key_rotary_pos_emb``; restated from
          the published 4.43.3 source) is injected before import (cache.py:4).
  shim 2  ``LlamaAttention.__init__`` no longer sets ``num_heads / num_key_value_heads / hidden_size``;
          they are back-filled from the config after construction (modules.py:67-69,99).

and three run-time repairs (monkey-patches, never edits) make KV-cached decode reachable:

  R1 (D1) the cache object is built with ``object.__new__`` + the six dicts of cache.py:14-19,
          because cache.py:12 (``super.__init__``) raises TypeError;
  R2 (D2) ``get_seq_length()`` called without an id (model.py:91) resolves to the generation in flight;
  R3 (D3) spans are built with local layer ids ``0..n-1`` so that cache.py:103-105 list indexing works.
"""
from __future__ import annotations

import os
import sys
import types
from typing import List, Optional

import torch

REFERENCE_ROOT = os.environ.get("DLI_REFERENCE_ROOT", "/root/reference")
_PKG = "distributed_llm_inference"
_loaded = None


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, _PKG, "models", "llama"))


def _install_sinkcache_stub():
    import transformers.cache_utils as cu

    if hasattr(cu, "SinkCache"):
        return

    class SinkCache:  # minimal stand-in for transformers 4.43.3 SinkCache (helpers only)
        def __init__(self, window_length: int, num_sink_tokens: int) -> None:
            self.window_length = window_length
            self.num_sink_tokens = num_sink_tokens

        @staticmethod
        def _rotate_half(x):
            x1 = x[..., : x.shape[-1] // 2]
            x2 = x[..., x.shape[-1] // 2:]
            return torch.cat((-x2, x1), dim=-1)

        def _apply_key_rotary_pos_emb(self, key_states, cos, sin):
            return (key_states * cos) + (self._rotate_half(key_states) * sin)

    cu.SinkCache = SinkCache


def load_reference():
    """Import the reference's llama modules from REFERENCE_ROOT without disturbing any module of the
    same top-level name that this repo provides (the drop-in alias package)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f"reference checkout not found at {REFERENCE_ROOT}")
    _install_sinkcache_stub()
    saved = {k: v for k, v in sys.modules.items() if k == _PKG or k.startswith(_PKG + ".")}
    for k in saved:
        del sys.modules[k]
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        import importlib

        model = importlib.import_module(_PKG + ".models.llama.model")
        modules = importlib.import_module(_PKG + ".models.llama.modules")
        cache = importlib.import_module(_PKG + ".models.llama.cache")
        assert model.__file__.startswith(REFERENCE_ROOT), model.__file__
    finally:
        sys.path.remove(REFERENCE_ROOT)
        for k in [k for k in sys.modules if k == _PKG or k.startswith(_PKG + ".")]:
            del sys.modules[k]
        sys.modules.update(saved)
    # transformers resolves sys.modules[cls.__module__] at construction: park the reference modules
    # under private names and point their classes there.
    for tag, mod in (("model", model), ("modules", modules), ("cache", cache)):
        alias = f"_dli_upstream_{tag}"
        sys.modules[alias] = mod
        for obj in vars(mod).values():
            if isinstance(obj, type) and getattr(obj, "__module__", "") == mod.__name__:
                obj.__module__ = alias
    _loaded = types.SimpleNamespace(model=model, modules=modules, cache=cache)
    return _loaded


def make_config(hidden_size, intermediate_size, num_attention_heads, num_key_value_heads,
                attn_implementation: str = "eager", rope_theta: float = 10000.0,
                max_position_embeddings: int = 8192):
    from transformers import LlamaConfig

    cfg = LlamaConfig(
        hidden_size=hidden_size, intermediate_size=intermediate_size,
        num_attention_heads=num_attention_heads, num_key_value_heads=num_key_value_heads,
        num_hidden_layers=1, vocab_size=32000, max_position_embeddings=max_position_embeddings,
        rope_parameters={"rope_type": "default", "rope_theta": rope_theta},
        attention_bias=False, mlp_bias=False, pretraining_tp=1,
    )
    cfg._attn_implementation = attn_implementation
    return cfg


def make_block(cfg, n_layers: int, rope: str = "literal", dtype=torch.float32):
    """Build a reference ``LlamaBlock`` with local layer ids (R3) and back-filled attributes (shim 2).
    ``rope="float"`` wraps ``rotary_emb.forward`` so that its first argument carries the activation
    dtype (what ``rotary_emb(hidden_states, position_ids)`` would pass) without touching the file."""
    ref = load_reference()
    impl = cfg._attn_implementation
    blk = ref.model.LlamaBlock(cfg, list(range(n_layers)))
    blk.config._attn_implementation = impl
    for layer in blk.layers:
        a = layer.self_attn
        a.num_heads = cfg.num_attention_heads
        a.num_key_value_heads = cfg.num_key_value_heads
        a.hidden_size = cfg.hidden_size
        a.config._attn_implementation = impl
    blk = blk.to(dtype).eval()
    if rope == "float":
        orig = blk.rotary_emb.forward

        def fwd(x, position_ids, _orig=orig, _dt=dtype):
            return _orig(torch.empty(0, dtype=_dt), position_ids)

        blk.rotary_emb.forward = fwd
    elif rope != "literal":
        raise ValueError(rope)
    return blk


def make_cache(window_length: int, num_sink_tokens: int):
    """R1 + R2."""
    ref = load_reference()
    base = ref.cache.PartialLlamaSinkCache

    class RepairedCache(base):
        def __init__(self, window_length, num_sink_tokens):  # R1: do not call the crashing __init__
            self.window_length = window_length
            self.num_sink_tokens = num_sink_tokens
            self.key_cache, self.value_cache = {}, {}
            self._cos_cache, self._sin_cache = {}, {}
            self._seen_tokens, self.cos_sin_rerotation_cache = {}, {}
            self._gid_in_flight: Optional[str] = None

        def get_seq_length(self, layer_idx=0, generation_id=None):  # R2
            if generation_id is None:
                generation_id = self._gid_in_flight
            return base.get_seq_length(self, layer_idx, generation_id)

    return RepairedCache(window_length, num_sink_tokens)


@torch.no_grad()
def block_forward(blk, generation_id: str, hidden, cache=None, cache_position=None,
                  output_hidden_states: bool = False, attention_mask=None, position_ids=None):
    if cache is not None:
        cache._gid_in_flight = generation_id
    return blk(generation_id, hidden, attention_mask=attention_mask, position_ids=position_ids,
               past_key_value=cache, output_hidden_states=output_hidden_states,
               cache_position=cache_position)


def load_layer_weights(blk, layers: List[dict]):
    """Copy a list of ``{name: tensor}`` dicts (reference state-dict names, per local layer) into blk."""
    sd = blk.state_dict()
    with torch.no_grad():
        for j, lw in enumerate(layers):
            for name, t in lw.items():
                sd[f"layers.{j}.{name}.weight"].copy_(t.to(sd[f"layers.{j}.{name}.weight"].dtype))
