"""CPU oracle for the span forward — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

This file is a plain-PyTorch fp32 restatement of the reference's hot path (the forward of a
contiguous span of Llama decoder blocks with a per-generation sink KV cache).  It exists to CHECK
the CUDA path; nothing under ``distributed_llm_inference_b200/`` imports it.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``
may import it.

Every function cites the reference file:line it follows (paths relative to /root/reference, the
upstream checkout of Dylan102938/distributed-llm-inference).  Arithmetic that the reference takes
from ``transformers==4.43.3`` (not vendored upstream; pinned in poetry.lock:1799-1800) is restated
from that version's published source (LlamaRMSNorm, LlamaMLP, LlamaRotaryEmbedding, rotate_half,
repeat_kv, SinkCache helpers).

Parity pinning: the reference ships NO tests, golden vectors or fixtures for this path
(tests/__init__.py and tests/server/__init__.py are empty), so the oracle is pinned against outputs
of the reference ITSELF: ``oracle/ref_shim.py`` executes the untouched reference files from
/root/reference, ``tests/golden/make_golden.py`` stores their outputs as fixtures under
``tests/golden/``, and ``tests/test_oracle_golden.py`` compares this restatement with them.

Crash repairs.  Three reference defects make KV-cached decode impossible as written:
the cache constructor raises (cache.py:12), ``get_seq_length()`` is called without the mandatory
generation id (model.py:91 vs cache.py:50-52), and global layer ids index a per-span list
(model.py:22 vs cache.py:103-105).  The oracle restates the REPAIRED behaviour (construct the six
dicts, resolve the missing id to the generation in flight, index by local position in the span) and
keeps every numerical quirk literally: RMSNorm eps 1e-6 (D7), post-attention norm of ``h + h``
(D6), cos/sin truncated to int64 in ``rope="literal"`` mode (D4), and no causal mask when the
attention implementation is ``sdpa`` (D5).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F


# ----------------------------------------------------------------------------------------------
# configuration
# ----------------------------------------------------------------------------------------------
@dataclass
class SpanShape:
    """Llama shape parameters used by the span (mirrors the LlamaConfig fields the reference reads:
    modules.py:119-125, transformers LlamaAttention.__init__, LlamaMLP.__init__)."""

    hidden_size: int
    intermediate_size: int
    num_attention_heads: int
    num_key_value_heads: int
    head_dim: int
    rope_theta: float = 10000.0
    rms_eps: float = 1e-6  # modules.py:124-125 -> LlamaRMSNorm(hidden_size) default eps (D7)


# ----------------------------------------------------------------------------------------------
# elementary ops
# ----------------------------------------------------------------------------------------------

# The seven linear layers of a decoder layer go through this name, so that the tests of the quantised modes can swap in
# the arithmetic of oracle/quant_oracle.py (int8 weights and/or activations); by default it is the reference's nn.Linear.
LINEAR = F.linear

def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    """transformers 4.43.3 ``LlamaRMSNorm.forward`` as called at modules.py:162,179."""
    dt = x.dtype
    x32 = x.to(torch.float32)
    var = x32.pow(2).mean(-1, keepdim=True)
    y = x32 * torch.rsqrt(var + eps)
    return weight * y.to(dt)


def rope_tables(
    position_ids: torch.Tensor, head_dim: int, theta: float, mode: str, act_dtype: torch.dtype
) -> Tuple[torch.Tensor, torch.Tensor]:
    """transformers 4.43.3 ``LlamaRotaryEmbedding.forward`` as called at model.py:55.

    ``position_ids`` is ``[Bp, L]`` int64.  The reference passes ``position_ids`` as BOTH arguments,
    so the tables are cast to int64 (truncation toward zero) — ``mode="literal"`` (D4).
    ``mode="float"`` is the intended behaviour: cast to the activation dtype.
    Returns cos, sin of shape ``[Bp, L, head_dim]``.
    """
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).float() / head_dim))
    inv_freq = inv_freq.to(position_ids.device)   # the oracle also runs in fp32 on a GPU for full-depth checks
    inv = inv_freq[None, :, None].float().expand(position_ids.shape[0], -1, 1)
    pos = position_ids[:, None, :].float()
    freqs = (inv @ pos).transpose(1, 2)
    emb = torch.cat((freqs, freqs), dim=-1)
    cos, sin = emb.cos(), emb.sin()
    if mode == "literal":
        return cos.to(torch.int64), sin.to(torch.int64)
    if mode == "float":
        return cos.to(act_dtype), sin.to(act_dtype)
    raise ValueError(f"unknown rope mode {mode!r}")


def rotate_half(x: torch.Tensor) -> torch.Tensor:
    """transformers ``rotate_half`` (imported at modules.py:6-10)."""
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


def apply_rotary_pos_emb(q, k, cos, sin):
    """modules.py:17-20 — note: no unsqueeze, cos ``[Bp, L, D]`` broadcasts against ``[B, H, L, D]``
    so only ``Bp == 1`` is meaningful (D9)."""
    return (q * cos) + (rotate_half(q) * sin), (k * cos) + (rotate_half(k) * sin)


def repeat_kv(x: torch.Tensor, n_rep: int) -> torch.Tensor:
    """transformers ``repeat_kv`` as called at modules.py:87-88 (repeat-interleave on the head dim)."""
    if n_rep == 1:
        return x
    b, kv, s, d = x.shape
    return x[:, :, None, :, :].expand(b, kv, n_rep, s, d).reshape(b, kv * n_rep, s, d)


# ----------------------------------------------------------------------------------------------
# KV cache with attention sinks (cache.py:7-135, repaired per D1-D3)
# ----------------------------------------------------------------------------------------------
class SinkCacheOracle:
    """Restatement of ``PartialLlamaSinkCache`` (cache.py:7-135)."""

    def __init__(self, window_length: int, num_sink_tokens: int):
        # cache.py:11-19 (the super().__init__ call of :12 is the D1 crash; state restated here)
        self.window_length = window_length
        self.num_sink_tokens = num_sink_tokens
        self.key_cache: Dict[str, List[torch.Tensor]] = {}
        self.value_cache: Dict[str, List[torch.Tensor]] = {}
        self._cos_cache: Dict[str, Optional[torch.Tensor]] = {}
        self._sin_cache: Dict[str, Optional[torch.Tensor]] = {}
        self._seen_tokens: Dict[str, int] = {}
        self.cos_sin_rerotation_cache: Dict[str, Dict[int, Tuple[torch.Tensor, torch.Tensor]]] = {}

    def _get_rerotation_cos_sin(self, key_states, cos, sin, generation_id):
        """cache.py:21-48."""
        if generation_id is None:
            raise ValueError("generation_id not provided")
        memo = self.cos_sin_rerotation_cache.get(generation_id)
        if memo is None:
            raise ValueError("cos_sin_rerotation_cache not found")
        n = key_states.shape[-2]
        if n not in memo:
            cos = cos.to(torch.float32)
            sin = sin.to(torch.float32)
            s = self.num_sink_tokens
            original_cos, shifted_cos = cos[s + n:], cos[s:-n]
            original_sin, shifted_sin = sin[s + n:], sin[s:-n]
            rc = original_cos * shifted_cos + original_sin * shifted_sin
            rs = -original_sin * shifted_cos + original_cos * shifted_sin
            memo[n] = (rc.to(key_states.dtype).unsqueeze(0), rs.to(key_states.dtype).unsqueeze(0))
        return memo[n]

    def get_seq_length(self, layer_idx: int = 0, generation_id: Optional[str] = None) -> int:
        """cache.py:50-62."""
        if not generation_id:
            raise ValueError("generation_id not provided")
        kc = self.key_cache.get(generation_id)
        if not kc or len(kc) <= layer_idx:
            return 0
        return kc[layer_idx].shape[-2]

    def update(self, key_states, value_states, layer_idx: int, cache_kwargs=None):
        """cache.py:64-135.  ``layer_idx`` is the LOCAL position in the span (D3 repair)."""
        if cache_kwargs is None or "generation_id" not in cache_kwargs:
            raise ValueError("generation_id not found in cache_kwargs")
        gid = cache_kwargs.get("generation_id")
        sin, cos = cache_kwargs.get("sin"), cache_kwargs.get("cos")
        if gid not in self.key_cache:  # cache.py:78-84
            self.key_cache[gid], self.value_cache[gid] = [], []
            self._cos_cache[gid] = self._sin_cache[gid] = None
            self._seen_tokens[gid] = 0
            self.cos_sin_rerotation_cache[gid] = {}
        n = key_states.shape[-2]
        if layer_idx == 0:  # cache.py:86-87
            self._seen_tokens[gid] += n
        using_rope = sin is not None and cos is not None
        if using_rope and layer_idx == 0:  # cache.py:91-101
            if cos.dim() == 2:
                self._cos_cache[gid], self._sin_cache[gid] = cos, sin
            elif self._cos_cache[gid] is None:
                self._cos_cache[gid], self._sin_cache[gid] = cos[0, ...], sin[0, ...]
            elif self._cos_cache[gid].shape[0] < self.window_length:
                self._cos_cache[gid] = torch.cat([self._cos_cache[gid], cos[0, ...]], dim=0)
                self._sin_cache[gid] = torch.cat([self._sin_cache[gid], sin[0, ...]], dim=0)
        kc, vc = self.key_cache[gid], self.value_cache[gid]
        if len(kc) <= layer_idx:  # cache.py:103-105: first chunk stored whole, even past the window (D14)
            kc.append(key_states)
            vc.append(value_states)
        elif n + self.get_seq_length(layer_idx, gid) < self.window_length:  # cache.py:107-109
            kc[layer_idx] = torch.cat([kc[layer_idx], key_states], dim=-2)
            vc[layer_idx] = torch.cat([vc[layer_idx], value_states], dim=-2)
        else:  # cache.py:111-133: [sinks | re-rotated tail | new]
            W, s = self.window_length, self.num_sink_tokens
            keep = kc[layer_idx][:, :, -W + s + n:]
            if using_rope:
                rc, rs = self._get_rerotation_cos_sin(
                    key_states, self._cos_cache[gid][:W], self._sin_cache[gid][:W], gid
                )
                keep = (keep * rc) + (rotate_half(keep) * rs)  # SinkCache._apply_key_rotary_pos_emb
            kc[layer_idx] = torch.cat([kc[layer_idx][:, :, :s], keep, key_states], dim=-2)
            vkeep = vc[layer_idx][:, :, -W + s + n:]
            vc[layer_idx] = torch.cat([vc[layer_idx][:, :, :s], vkeep, value_states], dim=-2)
        return kc[layer_idx], vc[layer_idx]


# ----------------------------------------------------------------------------------------------
# mask (model.py:78-144)
# ----------------------------------------------------------------------------------------------
def build_mask(
    attn_implementation: str,
    attention_mask: Optional[torch.Tensor],
    hidden: torch.Tensor,
    cache_position: torch.Tensor,
    past_seen: int,
) -> Optional[torch.Tensor]:
    """``LlamaBlock._update_causal_mask`` (model.py:78-144) for the cases the span supports:
    no padding mask (or an all-ones 2-D mask).  ``sdpa``/``flash_attention_2`` => ``None`` (D5,
    model.py:86-101); ``eager`` => additive 4-D causal mask (model.py:103-127)."""
    if attn_implementation == "flash_attention_2":
        if attention_mask is not None and bool((attention_mask == 0).any()):
            return attention_mask
        return None
    if attn_implementation == "sdpa":
        # AttentionMaskConverter._ignore_causal_mask_sdpa returns True when there is no padding to
        # mask (attention_mask None or all ones) and q_len == 1 or past == 0 ... the reference then
        # returns None and modules.py:92-94 adds nothing -> bidirectional prefill.
        if attention_mask is None or bool(torch.all(attention_mask == 1)):
            if hidden.shape[1] == 1 or past_seen == 0:
                return None
    dtype = hidden.dtype
    min_dtype = torch.finfo(dtype).min
    L = hidden.shape[1]
    target = attention_mask.shape[-1] if isinstance(attention_mask, torch.Tensor) else past_seen + L + 1
    m = torch.full((L, target), fill_value=min_dtype, dtype=dtype, device=hidden.device)
    if L != 1:
        m = torch.triu(m, diagonal=1)
    m = m * (torch.arange(target, device=hidden.device) > cache_position.reshape(-1, 1))
    m = m[None, None, :, :].expand(hidden.shape[0], 1, -1, -1)
    if attention_mask is not None:
        m = m.clone()
        ml = attention_mask.shape[-1]
        pad = (m[:, :, :, :ml] + attention_mask[:, None, None, :]) == 0
        m[:, :, :, :ml] = m[:, :, :, :ml].masked_fill(pad, min_dtype)
    return m


# ----------------------------------------------------------------------------------------------
# layer and span
# ----------------------------------------------------------------------------------------------
@dataclass
class LayerWeights:
    """Tensors named as in the reference state dict (modules.py:122-125 + HF LlamaAttention/LlamaMLP)."""

    q_proj: torch.Tensor
    k_proj: torch.Tensor
    v_proj: torch.Tensor
    o_proj: torch.Tensor
    gate_proj: torch.Tensor
    up_proj: torch.Tensor
    down_proj: torch.Tensor
    input_layernorm: torch.Tensor
    post_attention_layernorm: torch.Tensor

    @staticmethod
    def from_state_dict(sd: Dict[str, torch.Tensor], j: int, dtype=torch.float32) -> "LayerWeights":
        p = f"layers.{j}."
        g = lambda n: sd[p + n + ".weight"].detach().to(dtype).contiguous()
        return LayerWeights(
            g("self_attn.q_proj"), g("self_attn.k_proj"), g("self_attn.v_proj"), g("self_attn.o_proj"),
            g("mlp.gate_proj"), g("mlp.up_proj"), g("mlp.down_proj"),
            g("input_layernorm"), g("post_attention_layernorm"),
        )


def attention_forward(
    shape: SpanShape, w: LayerWeights, x: torch.Tensor, generation_id: str, cos, sin,
    mask: Optional[torch.Tensor], cache: Optional[SinkCacheOracle], local_layer_idx: int,
) -> torch.Tensor:
    """``OptimizedLlamaInferenceAttention.forward`` (modules.py:36-114), pretraining_tp == 1."""
    H, KV, hd = shape.num_attention_heads, shape.num_key_value_heads, shape.head_dim
    bsz, q_len, _ = x.shape
    q = LINEAR(x, w.q_proj).view(bsz, q_len, H, hd).transpose(1, 2)      # :61,67
    k = LINEAR(x, w.k_proj).view(bsz, q_len, KV, hd).transpose(1, 2)     # :62,68
    v = LINEAR(x, w.v_proj).view(bsz, q_len, KV, hd).transpose(1, 2)     # :63,69
    q, k = apply_rotary_pos_emb(q, k, cos, sin)                            # :73-76
    if cache is not None:                                                  # :78-85
        k, v = cache.update(k, v, layer_idx=local_layer_idx,
                            cache_kwargs={"cos": cos, "sin": sin, "generation_id": generation_id})
    k = repeat_kv(k, H // KV)                                              # :87-88
    v = repeat_kv(v, H // KV)
    s = torch.matmul(q, k.transpose(2, 3)) / math.sqrt(hd)                 # :90
    if mask is not None:                                                   # :92-94
        s = s + mask[:, :, :, : k.shape[-2]]
    p = F.softmax(s, dim=-1, dtype=torch.float32).to(q.dtype)              # :96
    o = torch.matmul(p, v)                                                 # :97
    o = o.transpose(1, 2).contiguous().reshape(bsz, q_len, -1)             # :105
    return LINEAR(o, w.o_proj)                                           # :112


def layer_forward(shape, w, x, generation_id, cos, sin, mask, cache, local_layer_idx):
    """``OptimizedLlamaDecoderLayer.forward`` (modules.py:146-184)."""
    residual = x
    h = rmsnorm(x, w.input_layernorm, shape.rms_eps)                                   # :162
    h = attention_forward(shape, w, h, generation_id, cos, sin, mask, cache, local_layer_idx)
    h = h + residual                                                                   # :173
    residual = h
    h = rmsnorm(h + residual, w.post_attention_layernorm, shape.rms_eps)               # :179 (D6: norm of 2h)
    h = LINEAR(F.silu(LINEAR(h, w.gate_proj)) * LINEAR(h, w.up_proj), w.down_proj)  # :181 LlamaMLP
    return h + residual                                                                # :182


@dataclass
class OracleSpan:
    """``LlamaBlock`` (model.py:16-76) over ``layers`` with repaired cache plumbing."""

    shape: SpanShape
    layers: List[LayerWeights]
    attn_implementation: str = "eager"   # "eager" => causal; "sdpa" => reference default, unmasked prefill (D5)
    rope: str = "literal"                # "literal" => int64 tables (D4); "float" => intended

    def forward(
        self,
        generation_id: str,
        hidden_states: torch.Tensor,
        attention_mask: Optional[torch.Tensor] = None,
        position_ids: Optional[torch.Tensor] = None,
        past_key_value: Optional[SinkCacheOracle] = None,
        output_hidden_states: bool = False,
        cache_position: Optional[torch.Tensor] = None,
    ):
        if cache_position is None:                                                     # model.py:39-42
            cache_position = torch.arange(0, hidden_states.shape[1], device=hidden_states.device)
        past_seen = 0
        if past_key_value is not None:                                                 # model.py:91 (D2 repaired)
            past_seen = past_key_value.get_seq_length(0, generation_id)
        mask = build_mask(self.attn_implementation, attention_mask, hidden_states, cache_position, past_seen)
        if position_ids is None:                                                       # model.py:52-53
            position_ids = cache_position.unsqueeze(0)
        cos, sin = rope_tables(position_ids, self.shape.head_dim, self.shape.rope_theta,
                               self.rope, hidden_states.dtype)                         # model.py:55
        all_h = () if output_hidden_states else None
        h = hidden_states
        for j, w in enumerate(self.layers):                                            # model.py:59-71
            if output_hidden_states:
                all_h += (h,)
            h = layer_forward(self.shape, w, h, generation_id, cos, sin, mask, past_key_value, j)
        if output_hidden_states:
            all_h += (h,)
        return tuple(v for v in (h, all_h) if v is not None)                           # model.py:76

    __call__ = forward


# ----------------------------------------------------------------------------------------------
# harness client (absent upstream: model.py:19-23 keeps vocab_size/padding_idx but builds no
# embedding, final norm or head).  embed -> spans -> RMSNorm -> lm_head -> argmax.
# ----------------------------------------------------------------------------------------------
@dataclass
class OracleClient:
    embed: torch.Tensor          # [vocab, h]
    final_norm: torch.Tensor     # [h]
    lm_head: torch.Tensor        # [vocab, h]
    spans: Sequence[OracleSpan]
    caches: Sequence[SinkCacheOracle]
    norm_eps: float = 1e-6
    logits_log: List[torch.Tensor] = field(default_factory=list)

    def step(self, generation_id: str, token_ids: torch.Tensor, start_pos: int) -> torch.Tensor:
        """Run ``token_ids`` ``[B, L]`` at absolute positions ``start_pos..`` through every span;
        return last-position logits ``[B, vocab]`` (fp32)."""
        h = F.embedding(token_ids, self.embed)
        cp = torch.arange(start_pos, start_pos + token_ids.shape[1])
        for span, cache in zip(self.spans, self.caches):
            h = span.forward(generation_id, h, past_key_value=cache, cache_position=cp)[0]
        h = rmsnorm(h[:, -1:, :], self.final_norm, self.norm_eps)
        return F.linear(h[:, 0, :], self.lm_head)

    def greedy(self, generation_id: str, prompt: torch.Tensor, n_new: int,
               forced: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, List[torch.Tensor]]:
        """Greedy decode.  ``forced`` ``[B, n_new]`` teacher-forces the fed-back tokens (the argmax is
        still recorded) so two implementations stay comparable after a near-tie."""
        logits = self.step(generation_id, prompt, 0)
        out, logs = [], []
        pos = prompt.shape[1]
        for i in range(n_new):
            nxt = logits.argmax(-1)
            out.append(nxt)
            logs.append(logits)
            if i == n_new - 1:
                break
            feed = forced[:, i] if forced is not None else nxt
            logits = self.step(generation_id, feed[:, None], pos)
            pos += 1
        return torch.stack(out, dim=1), logs
