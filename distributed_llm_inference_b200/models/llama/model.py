"""``LlamaBlock`` — a contiguous span of Llama decoder layers executed by one C call into libdli_b200.

Drop-in for distributed_llm_inference/models/llama/model.py:16-144: same constructor
``LlamaBlock(config, layer_ids)`` (model.py:17), same ``forward`` signature and return tuple
(model.py:25-34,76), same parameter names.  What differs is underneath: the layer loop of model.py:59-71
runs as ``dli_span_forward`` (hand-written sm_100a kernels behind a C ABI), the KV cache is paged
(``PartialLlamaSinkCache`` in this package), the additive mask of model.py:78-144 collapses to a ``causal``
flag, and cos/sin (model.py:55) are gathered from a table so that both the reference's literal behaviour
(int64-truncated tables, SURVEY D4: ``rope="literal"``) and the intended one (``rope="float"``) are host-side
choices.

There is NO CPU or PyTorch fallback: inputs must be bf16 CUDA tensors on a B200 and the shared library
must be present; anything else raises.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional

import torch
from torch import nn

from .cache import PartialLlamaSinkCache
from .modules import OptimizedLlamaDecoderLayer

_NOCACHE_ID = "__dli_no_cache__"


def rope_table(head_dim: int, theta: float, max_pos: int, mode: str, device) -> torch.Tensor:
    """``[2, max_pos, head_dim]`` fp32 (cos, sin) following transformers ``LlamaRotaryEmbedding.forward`` as the
    reference calls it (model.py:55).  ``literal``: the reference passes ``position_ids`` as the dtype carrier,
    so the tables are truncated toward zero to int64 — stored here as the same integers in fp32."""
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).float() / head_dim))
    pos = torch.arange(max_pos, dtype=torch.float32)
    freqs = (inv_freq[None, :, None].float() @ pos[None, None, :].float()).transpose(1, 2)[0]   # [max_pos, hd/2]
    emb = torch.cat((freqs, freqs), dim=-1)
    cos, sin = emb.cos(), emb.sin()
    if mode == "literal":
        cos, sin = cos.to(torch.int64).float(), sin.to(torch.int64).float()
    elif mode != "float":
        raise ValueError(f"rope must be 'literal' or 'float', got {mode!r}")
    return torch.stack([cos, sin]).contiguous().to(device)


@dataclass
class StageIO:
    """Inputs, outputs and hand-off flags of a pipeline stage step (``dli_stage_io`` in include/dli_b200.h), as tensors:
    views of this GPU's memory, or of a peer GPU's symmetric-memory buffer for ``hidden_out`` / ``tokens_out`` /
    ``signal_flag``.  ``parallel/pipeline.py`` builds these; ``GraphedDecode`` keeps them alive."""
    hidden_in: Optional[torch.Tensor] = None       # bf16 [B, 1, h]
    token_ids: Optional[torch.Tensor] = None       # int32 [B] (with embed_table: stage 0 of a token ring)
    embed_table: Optional[torch.Tensor] = None     # bf16 [V, h]
    hidden_out: Optional[torch.Tensor] = None      # bf16 [B, 1, h]
    final_norm: Optional[torch.Tensor] = None      # bf16 [h]
    lm_head: Optional[torch.Tensor] = None         # bf16 [V, h]
    logits: Optional[torch.Tensor] = None          # fp32 [B, V]
    tokens_out: Optional[torch.Tensor] = None      # int32 [B]
    eps_final: float = 1e-6
    wait_flag: Optional[torch.Tensor] = None       # int32 view, word 0 is the flag
    signal_flag: Optional[torch.Tensor] = None
    state: Optional[torch.Tensor] = None           # int32 [4]
    wait_lag: int = 0
    chain: int = 0
    timeout_ms: int = 10000       # a rank held up by the host for seconds must not turn into a spurious time-out

    def to_ffi(self, ffi):
        p = lambda t: None if t is None else t.data_ptr()
        for name, t, dt in (("hidden_in", self.hidden_in, torch.bfloat16), ("hidden_out", self.hidden_out, torch.bfloat16),
                            ("embed_table", self.embed_table, torch.bfloat16), ("lm_head", self.lm_head, torch.bfloat16),
                            ("final_norm", self.final_norm, torch.bfloat16), ("logits", self.logits, torch.float32),
                            ("token_ids", self.token_ids, torch.int32), ("tokens_out", self.tokens_out, torch.int32),
                            ("wait_flag", self.wait_flag, torch.int32), ("signal_flag", self.signal_flag, torch.int32),
                            ("state", self.state, torch.int32)):
            if t is not None and (t.dtype != dt or not t.is_contiguous()):
                raise TypeError(f"StageIO.{name} must be a contiguous {dt} tensor")
        return ffi.StageIo(p(self.hidden_in), p(self.token_ids), p(self.embed_table), p(self.hidden_out), p(self.final_norm),
                           p(self.lm_head), p(self.logits), p(self.tokens_out),
                           0 if self.lm_head is None else self.lm_head.shape[0], self.eps_final,
                           p(self.wait_flag), p(self.signal_flag), p(self.state), self.wait_lag, self.chain, self.timeout_ms)


class LlamaBlock(nn.Module):
    def __init__(self, config, layer_ids: List[int], *, rope: str = "literal", dtype=torch.bfloat16):
        """``rope="literal"`` reproduces model.py:55 as written (int64 tables); ``rope="float"`` is what the
        reference evidently intended."""
        super().__init__()
        if dtype != torch.bfloat16:
            raise NotImplementedError("the B200 span kernels are bf16 (the north-star dtype)")
        self.config = config
        self.padding_idx = getattr(config, "pad_token_id", None)       # model.py:19
        self.vocab_size = getattr(config, "vocab_size", None)          # model.py:20
        self.layer_ids = list(layer_ids)
        self.rope = rope
        # layers keep their GLOBAL id as in model.py:22; the cache is indexed by local position
        self.layers = nn.ModuleList([OptimizedLlamaDecoderLayer(config, idx, dtype=dtype) for idx in self.layer_ids])
        self.head_dim = getattr(config, "head_dim", None) or config.hidden_size // config.num_attention_heads
        rp = getattr(config, "rope_parameters", None) or {}
        self.rope_theta = float(rp.get("rope_theta", getattr(config, "rope_theta", 10000.0)))
        if rp.get("rope_type", "default") != "default":
            raise NotImplementedError("only rope_type='default' is implemented")
        # rows of the cos/sin table; grown on demand (the reference computes cos/sin for arbitrary positions, model.py:55,
        # so its sink cache keeps generating past max_position_embeddings)
        self.max_positions = int(getattr(config, "max_position_embeddings", 4096))
        self._span = None           # dli_span* (ctypes void pointer)
        self._bound_ptrs = None
        self._table: Optional[torch.Tensor] = None
        self._ws: Optional[torch.Tensor] = None
        self._ws_key = (0, 0, 0)
        self._scratch_cache: Optional[PartialLlamaSinkCache] = None
        self._dq: Optional[torch.Tensor] = None
        self._a8_threshold: Optional[float] = None     # quantize_w8(activations="int8"): LLM.int8 threshold, None = bf16 activations
        self._a8_applied = None
        self._a8: Optional[torch.Tensor] = None
        self.last_launch_count = 0

    # ------------------------------------------------------------------ native span object
    def _ffi(self):
        from distributed_llm_inference_b200 import _ffi

        return _ffi

    def _ensure_span(self, device):
        ffi = self._ffi()
        for layer in self.layers:
            layer.fuse()
        w8 = [layer.w8 is not None for layer in self.layers]
        if any(w8) and not all(w8):
            raise ValueError("either every layer of a span is quantised (quantize_w8) or none")
        if all(w8):
            ptrs = tuple(t.data_ptr() for layer in self.layers
                         for t in (*layer.w8["qkv"], *layer.w8["o"], *layer.w8["gu"], *layer.w8["down"],
                                   layer.input_layernorm.weight, layer.post_attention_layernorm.weight))
        else:
            ptrs = tuple(p.data_ptr() for layer in self.layers for p in
                         (layer._wqkv, layer.self_attn.o_proj.weight, layer._wgu, layer.mlp.down_proj.weight,
                          layer.input_layernorm.weight, layer.post_attention_layernorm.weight))
        if self._span is not None and ptrs == self._bound_ptrs:
            self._sync_act_i8(device)
            return
        cfg = self.config
        if self._span is None:
            desc = ffi.SpanDesc(cfg.hidden_size, cfg.intermediate_size, cfg.num_attention_heads,
                                cfg.num_key_value_heads, self.head_dim, len(self.layers), 0,
                                1e-6, 1e-6, 2.0)   # modules.py:124-125 (eps), modules.py:179 (norm of h + h)
            out = C.c_void_p()
            ffi.check(ffi.lib.dli_span_create(C.byref(desc), C.byref(out)))
            self._span = out
        per = len(ptrs) // len(self.layers)
        for j, layer in enumerate(self.layers):
            for p in (layer.input_layernorm.weight, layer.post_attention_layernorm.weight):
                if p.device != device or p.dtype != torch.bfloat16:
                    raise ValueError(f"layer {j}: parameters must be bf16 on {device} (got {p.dtype} on {p.device})")
            if all(w8):
                w = ffi.LayerWeightsW8(*ptrs[per * j: per * (j + 1)])
                ffi.check(ffi.lib.dli_span_bind_weights_w8(self._span, j, C.byref(w)))
            else:
                for p in layer.parameters():
                    if p.device != device or p.dtype != torch.bfloat16:
                        raise ValueError(f"layer {j}: parameters must be bf16 on {device} (got {p.dtype} on {p.device})")
                w = ffi.LayerWeights(*ptrs[per * j: per * (j + 1)])
                ffi.check(ffi.lib.dli_span_bind_weights(self._span, j, C.byref(w)))
        if all(w8):   # prefill of a quantised span widens one matrix at a time into this scratch
            nbytes = ffi.lib.dli_span_dequant_scratch_bytes(self._span)
            if self._dq is None or self._dq.numel() < nbytes or self._dq.device != device:
                self._dq = torch.empty(nbytes, dtype=torch.uint8, device=device)
            ffi.check(ffi.lib.dli_span_set_dequant_scratch(self._span, self._dq.data_ptr(), self._dq.numel()))
        self._bound_ptrs = ptrs
        self._a8_applied = None
        self._sync_act_i8(device)

    def _sync_act_i8(self, device):
        """Hands the span its int8-activation scratch (or takes it away) when the requested mode changed."""
        want = self._a8_threshold if all(layer.w8 is not None for layer in self.layers) else None
        if want == self._a8_applied:
            return
        ffi = self._ffi()
        if want is None:
            ffi.check(ffi.lib.dli_span_set_act_i8(self._span, None, 0, 0.0))
        else:
            nbytes = ffi.lib.dli_span_act_i8_scratch_bytes(self._span)
            if self._a8 is None or self._a8.numel() < nbytes or self._a8.device != device:
                self._a8 = torch.empty(nbytes, dtype=torch.uint8, device=device)
            ffi.check(ffi.lib.dli_span_set_act_i8(self._span, self._a8.data_ptr(), self._a8.numel(), float(want)))
        self._a8_applied = want

    def quantize_w8(self, activations: str = "bf16", threshold: float = 5.0) -> "LlamaBlock":
        """Row-wise int8 weights for every layer of the span (see ``OptimizedLlamaDecoderLayer.quantize_w8``).
        ``activations="int8"``: decode steps (chunks of <= 128 tokens) also quantise the ACTIVATIONS per token row and
        multiply int8 x int8 on the tensor cores, entries with ``|x| >= threshold`` kept in 16 bits -- the arithmetic of
        the reference's ``Linear8bitLt(threshold=...)`` (utils/model.py:93-123); ``"bf16"`` keeps 16-bit activations.
        Decode graphs captured before a change of mode keep the mode they were captured in: capture them again."""
        if activations not in ("bf16", "int8"):
            raise ValueError("activations must be 'bf16' or 'int8'")
        for layer in self.layers:
            if layer.w8 is None:
                layer.quantize_w8()
        self._a8_threshold = float(threshold) if activations == "int8" else None
        return self

    def __del__(self):
        try:
            if self._span is not None:
                self._ffi().lib.dli_span_destroy(self._span)
        except Exception:
            pass

    def _rope_table(self, device, min_rows: int = 0) -> torch.Tensor:
        """``[2, rows, hd]`` fp32 cos/sin table on ``device`` with at least ``min_rows`` rows (doubling growth; a captured
        decode graph holds pointers into the table it was captured with, so ``GraphedDecode`` keeps its own reference)."""
        t = self._table
        if t is None or t.device != device or t.shape[1] < min_rows:
            rows = self.max_positions
            while rows < min_rows:
                rows *= 2
            self.max_positions = rows
            self._table = t = rope_table(self.head_dim, self.rope_theta, rows, self.rope, device)
        return t

    def _workspace(self, tokens: int, batch: int, max_ctx: int, device) -> torch.Tensor:
        key = self._ws_key
        if self._ws is None or tokens > key[0] or batch > key[1] or max_ctx > key[2] or self._ws.device != device:
            key = (max(tokens, key[0]), max(batch, key[1]), max(-(-max_ctx // 1024) * 1024, key[2]))
            nbytes = self._ffi().lib.dli_span_workspace_bytes(self._span, *key)
            self._ws = torch.zeros(nbytes, dtype=torch.uint8, device=device)   # zero-filled once (C ABI contract)
            self._ws_key = key
        return self._ws

    # ------------------------------------------------------------------ whole-step CUDA graph
    def graphed_decode(self, generation_id: str, cache: PartialLlamaSinkCache, max_ctx: int,
                       x: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                       stage: Optional[StageIO] = None, streaming_steps: int = 0) -> "GraphedDecode":
        """Capture ONE decode step (q_len = 1) of this span for a live generation into a CUDA graph.

        The reference hides decode launch latency with three small per-layer graphs (utils/cuda.py:6-77 used at
        modules.py:28-34,130-144); here the whole stage step — every kernel of every layer, nothing else — is one
        graph (``dli_span_decode_step``).  Everything that changes between steps (sequence lengths, block table,
        position) lives in device memory: the kernels read the step's cos/sin row at the device-resident position and
        the span's last kernel advances position and lengths, so the graph is replayed unchanged until ``max_ctx``
        tokens.
        ``x`` / ``out`` (optional, bf16 ``[B, 1, h]``) replace the graph's own static input / output buffers; ``out`` may
        be a peer GPU's mapped memory, in which case the span's last kernel stores the hand-off tensor straight into
        the next pipeline stage's input over NVLink.
        ``stage`` (instead of ``x`` / ``out``): the step is a pipeline stage step (``dli_stage_decode_step``) whose entry
        kernel waits for the producer's flag (or gathers embedding rows by token id) and whose last kernel stores the
        result into the consumer's memory and releases its flag -- see ``StageIO``.
        ``streaming_steps > 0`` (a generation whose sink window is FULL, ``length == window_length``): the captured step is
        the steady state of attention-sink streaming -- evict one token and re-rotate the kept keys (cache.py:111-133),
        then the decode step -- replayable for that many steps; ``max_ctx`` is ignored (the context stays at the window).
        """
        return GraphedDecode(self, cache, generation_id, max_ctx, x=x, out=out, stage=stage, streaming_steps=streaming_steps)

    # ------------------------------------------------------------------ instrumentation (bench.py)
    def profile(self, enable: bool = True) -> None:
        """Record a CUDA event at every kernel-class boundary of the following forwards (dli_span_profile)."""
        ffi = self._ffi()
        if self._span is None:
            raise RuntimeError("run one forward before enabling profiling")
        ffi.check(ffi.lib.dli_span_profile(self._span, int(enable)))

    def profile_read(self):
        """-> {class: (milliseconds, intervals)} accumulated since the last read."""
        ffi = self._ffi()
        ms = (C.c_float * len(ffi.PROF_CLASSES))()
        cnt = (C.c_int * len(ffi.PROF_CLASSES))()
        ffi.check(ffi.lib.dli_span_profile_read(self._span, ms, cnt))
        return {name: (float(ms[i]), int(cnt[i])) for i, name in enumerate(ffi.PROF_CLASSES)}

    # ------------------------------------------------------------------ forward
    def _mask_plan(self, attention_mask, q_len: int, past: int, batch: int, device):
        """model.py:78-144 reduced to ``(causal, key_valid)``: one bit plus, for a 2-D mask that really contains zeros, a
        ``[B, past + q_len]`` uint8 validity map of the keys (model.py:128-135).  An all-ones mask is the same as none."""
        impl = getattr(self.config, "_attn_implementation", None) or "sdpa"
        key_valid = None
        if attention_mask is not None:
            if attention_mask.dim() != 2:
                raise NotImplementedError("custom 4-D attention masks are not supported by the B200 span")
            if not bool(torch.all(attention_mask == 1)):            # one host read per call that passes a mask
                if impl == "flash_attention_2":
                    raise NotImplementedError("flash_attention_2 with a padding mask: the reference indexes the 2-D mask with "
                                              "four indices and raises (model.py:86-89 -> modules.py:92-94)")
                if attention_mask.shape[0] != batch or attention_mask.shape[1] < past + q_len:
                    raise ValueError(f"attention_mask must be [{batch}, >= {past + q_len}], got {tuple(attention_mask.shape)}")
                key_valid = (attention_mask[:, :past + q_len] != 0).to(device=device, dtype=torch.uint8).contiguous()
        if impl == "flash_attention_2":
            return 0, None                             # model.py:86-89: mask None -> modules.py:92-94 adds nothing
        if impl == "sdpa" and key_valid is None and (q_len == 1 or past == 0):
            return 0, None                             # model.py:94-101 (_ignore_causal_mask_sdpa) -> None
        return 1, key_valid                            # model.py:103-135

    def _causal_flag(self, attention_mask, q_len: int, past: int) -> int:
        impl = getattr(self.config, "_attn_implementation", None) or "sdpa"
        if impl == "flash_attention_2" or (impl == "sdpa" and attention_mask is None and (q_len == 1 or past == 0)):
            return 0
        return 1

    @torch.no_grad()
    def forward(
        self,
        generation_id: str,
        hidden_states: torch.Tensor,
        attention_mask: Optional[torch.Tensor] = None,
        position_ids: Optional[torch.LongTensor] = None,
        past_key_value: Optional[PartialLlamaSinkCache] = None,
        output_hidden_states: Optional[bool] = None,
        cache_position: Optional[torch.LongTensor] = None,
    ):
        ffi = self._ffi()
        if not hidden_states.is_cuda:
            raise RuntimeError("LlamaBlock (B200) needs CUDA tensors: there is no CPU fallback for the span forward")
        if hidden_states.dtype != torch.bfloat16 or hidden_states.dim() != 3:
            raise TypeError(f"hidden_states must be bf16 [B, L, h], got {hidden_states.dtype} {tuple(hidden_states.shape)}")
        output_hidden_states = (output_hidden_states if output_hidden_states is not None
                                else bool(getattr(self.config, "output_hidden_states", False)))   # model.py:35-37
        device = hidden_states.device
        B, L, h = hidden_states.shape
        if h != self.config.hidden_size:
            raise ValueError(f"hidden size {h} != config.hidden_size {self.config.hidden_size}")
        with torch.cuda.device(device):
            self._ensure_span(device)

        cache = past_key_value
        gid = generation_id
        if cache is None:                               # cache-less call: attention over the chunk only
            sc = self._scratch_cache
            need = B * (-(-L // 16)) * 16
            if sc is None or sc.n_pages * sc.page_tokens < need or (sc._k is not None and sc._k.device != device):
                sc = self._scratch_cache = PartialLlamaSinkCache(1 << 30, 0, page_tokens=16, capacity_tokens=need)
            cache, gid = sc, _NOCACHE_ID
            cache.close(gid)
        elif not gid:
            raise ValueError("generation_id not provided")   # cache.py:51-52 via model.py:91
        cache.bind(len(self.layers), self.config.num_key_value_heads, self.head_dim, device)
        gen = cache.generation(gid)
        past = gen.length if gen is not None else 0

        cache_position_given, position_ids_given = cache_position is not None, position_ids is not None
        if cache_position is None:                      # model.py:39-42
            if past != 0 and self._causal_flag(attention_mask, L, past):
                raise ValueError("cache_position must be given when continuing a generation (the reference's default "
                                 "arange(q_len) would mask by relative index, model.py:39-42,110)")
            cache_position = torch.arange(0, L, device=device)
        if position_ids is None:                        # model.py:52-53
            position_ids = cache_position.unsqueeze(0)
        if position_ids.dim() != 2 or position_ids.shape[0] != 1 or position_ids.shape[1] != L:
            raise ValueError("position_ids must be [1, q_len]: the reference's apply_rotary_pos_emb broadcasts one "
                             "position vector over the batch (modules.py:17-20)")
        # positions of a generation are contiguous (the reference's sink cache assumes the same for its cos/sin history,
        # cache.py:91-101), so the host knows an upper bound without reading the device tensor: the table is grown to
        # cover it and the gather index is clamped, so an out-of-range id can never fault on the device
        explicit_pos = cache_position_given or position_ids_given
        first_call = gen is None or (gen.length == 0 and gen.seen_tokens == 0)
        if first_call:
            first_pos = int(position_ids[0, 0].item()) if explicit_pos else 0   # once per generation
        else:
            first_pos = gen.first_pos
        seen = 0 if gen is None else gen.seen_tokens
        table = self._rope_table(device, first_pos + seen + L)
        idx = position_ids[0].to(device).clamp(0, table.shape[1] - 1)
        cs = torch.index_select(table, 1, idx)                                      # [2, L, hd] fp32 (model.py:55)

        if gen is not None and cache.needs_eviction(gid, L):   # cache.py:107-133: the window is full
            p0, W = gen.first_pos, cache.window_length
            table = self._rope_table(device, p0 + W)
            cache.evict(gid, L, table[0, p0:p0 + W], table[1, p0:p0 + W])
            past = gen.length
        gen = cache.reserve(gid, B, past + L)
        if first_call:
            gen.first_pos = first_pos
        max_ctx = past + L
        causal, key_valid = self._mask_plan(attention_mask, L, past, B, device)
        # The reference's causal rule compares a query's ABSOLUTE cache_position with a key's SLOT in the cache
        # (model.py:110).  The two coincide -- and the kernels' causal flag applies -- as long as the generation started at
        # position 0 and nothing was evicted; otherwise (a chunk that crosses the sink window, cache.py:111-133 with
        # n > 1, or a generation that starts at a later position) the rule becomes a per-query key limit.
        key_limit = None
        if causal and L > 1 and first_pos + seen != past:
            key_limit = (cache_position.to(device) + 1).clamp(max=max_ctx).to(torch.int32).contiguous()
        x = hidden_states.contiguous()
        out = torch.empty_like(x)
        all_h = torch.empty(len(self.layers) + 1, B, L, h, dtype=x.dtype, device=device) if output_hidden_states else None
        kv = cache.kv_desc(gen, max_ctx)
        with torch.cuda.device(device):                  # kernels launch on the tensors' device, whatever is current
            ws = self._workspace(B * L, B, max_ctx, device)
            try:
                if key_valid is not None or key_limit is not None:
                    am = ffi.AttnMask(None if key_limit is None else key_limit.data_ptr(),
                                      None if key_valid is None else key_valid.data_ptr(),
                                      0 if key_valid is None else key_valid.shape[1])
                    ffi.check(ffi.lib.dli_span_forward_masked(
                        self._span, C.byref(kv), B, L, causal, C.byref(am), x.data_ptr(), out.data_ptr(), cs[0].data_ptr(),
                        cs[1].data_ptr(), all_h.data_ptr() if all_h is not None else None, ws.data_ptr(), ws.numel(),
                        ffi.current_stream(device)))
                else:
                    ffi.check(ffi.lib.dli_span_forward(
                        self._span, C.byref(kv), B, L, causal, x.data_ptr(), out.data_ptr(), cs[0].data_ptr(), cs[1].data_ptr(),
                        all_h.data_ptr() if all_h is not None else None, ws.data_ptr(), ws.numel(), ffi.current_stream(device)))
            except ffi.DliError:
                self._ws = None                          # a chain that stopped half-way leaves partial sums behind: the
                raise                                    # next call gets a freshly zeroed workspace (include/dli_b200.h)
        self.last_launch_count = ffi.lib.dli_last_launch_count()
        cache.advance(gid, L)
        if past_key_value is None:
            cache.close(gid)
        res = (out,)
        if all_h is not None:
            res = res + (tuple(all_h[i] for i in range(all_h.shape[0])),)
        return res                                       # model.py:76


class GraphedDecode:
    """A captured decode step: ``out = step(x)`` appends one token per sequence and returns the span output
    (a static buffer that the next ``step`` overwrites)."""

    def __init__(self, block: LlamaBlock, cache: PartialLlamaSinkCache, generation_id: str, max_ctx: int,
                 x: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None, stage: Optional[StageIO] = None,
                 streaming_steps: int = 0):
        gen = cache.generation(generation_id)
        if gen is None or gen.length == 0:
            raise ValueError("graphed_decode needs a generation that has been prefilled through this span")
        ffi = block._ffi()
        device = gen.seq_lens.device
        B, h = gen.batch, block.config.hidden_size
        W = cache.window_length
        self.streaming = streaming_steps > 0
        self.steps_left = streaming_steps
        if self.streaming:
            if gen.length != W or stage is not None:
                raise ValueError("a streaming graph is captured on a generation whose window is full "
                                 f"(length {gen.length}, window {W}) and outside a pipeline stage")
            if W - cache.num_sink_tokens - 1 <= 0:
                raise ValueError("window too small for its sink tokens")
            max_ctx = W
        self.block, self.cache, self.gid, self.max_ctx = block, cache, generation_id, max_ctx
        if not self.streaming and max_ctx >= W:
            raise ValueError("graphed decode cannot cross window_length; once the window is full capture the steady "
                             "state with streaming_steps")
        cache.reserve(generation_id, B, max_ctx)                    # block table fixed for the graph's lifetime
        with torch.cuda.device(device):
            block._ensure_span(device)
            ws = block._workspace(B, B, max_ctx, device)
        table = block._rope_table(device, gen.first_pos + gen.seen_tokens + max(max_ctx - gen.length, streaming_steps) + 1)
        for t in (x, out):
            if t is not None and (tuple(t.shape) != (B, 1, h) or t.dtype != torch.bfloat16 or not t.is_contiguous()):
                raise ValueError(f"graphed_decode buffers must be contiguous bf16 [{B}, 1, {h}]")
        self.x = x if x is not None else torch.zeros(B, 1, h, dtype=torch.bfloat16, device=device)
        self.out = out if out is not None else torch.empty(B, 1, h, dtype=torch.bfloat16, device=device)
        self.pos = torch.tensor([gen.first_pos + gen.seen_tokens], dtype=torch.long, device=device)
        kv = cache.kv_desc(gen, max_ctx)

        self.device = device
        self.stage = stage
        io = stage.to_ffi(ffi) if stage is not None else None

        def launch_stage():
            try:
                ffi.check(ffi.lib.dli_stage_decode_step(block._span, C.byref(kv), B, C.byref(io), table[0].data_ptr(),
                                                        table[1].data_ptr(), self.pos.data_ptr(), 1, ws.data_ptr(),
                                                        ws.numel(), ffi.current_stream(device)))
            except ffi.DliError:
                block._ws = None
                raise
            return table, ws, stage

        def launch_streaming():
            # steady state of the sink window, all constants: evict one token for every layer (keys re-rotated back by one
            # position with the generation's cos/sin history), lengths W -> W - 1, then the step appends and restores W
            p0 = gen.first_pos
            n_layers, KV, hd = cache._shape
            ch, sh = table[0, p0:p0 + W], table[1, p0:p0 + W]
            kvw = cache.kv_desc(gen, W)
            st = ffi.current_stream(device)
            try:
                ffi.check(ffi.lib.dli_kv_evict(C.byref(kvw), n_layers, B, KV, hd, W, W, cache.num_sink_tokens, 1, ch.data_ptr(),
                                               sh.data_ptr(), st))
                gen.seq_lens.fill_(W - 1)
                ffi.check(ffi.lib.dli_span_decode_step(block._span, C.byref(kv), B, self.x.data_ptr(), self.out.data_ptr(),
                                                       table[0].data_ptr(), table[1].data_ptr(), self.pos.data_ptr(), 1,
                                                       ws.data_ptr(), ws.numel(), st))
            except ffi.DliError:
                block._ws = None
                raise
            return table, ws, ch, sh

        def launch():
            # position and lengths live on the device: the kernels index the cos/sin tables by *pos and the span's last
            # kernel advances pos and seq_lens, so a replayed step is nothing but the span's own PDL-chained kernels
            try:
                ffi.check(ffi.lib.dli_span_decode_step(block._span, C.byref(kv), B, self.x.data_ptr(), self.out.data_ptr(),
                                                       table[0].data_ptr(), table[1].data_ptr(), self.pos.data_ptr(), 1,
                                                       ws.data_ptr(), ws.numel(), ffi.current_stream(device)))
            except ffi.DliError:
                block._ws = None
                raise
            return table, ws

        # run the decode-path kernels once outside capture (one-time function attributes, descriptor caches) on a
        # throw-away generation; the capture itself launches nothing, so the live generation is untouched
        warm = "__dli_graph_warmup__"
        block(warm, self.x, past_key_value=cache, cache_position=torch.zeros(1, dtype=torch.long, device=device))
        cache.close(warm)
        if self.streaming:
            launch = launch_streaming             # (the eviction kernel needs no one-time set-up)
        if stage is not None:
            launch = launch_stage
            if stage.lm_head is not None:     # the head's GEMM instantiation, outside capture as well
                V = stage.lm_head.shape[0]
                with torch.cuda.device(device):
                    ffi.check(ffi.lib.dli_lm_head_argmax(
                        self.x.data_ptr(), 0, stage.final_norm.data_ptr(), stage.eps_final, stage.lm_head.data_ptr(), V, h, B,
                        torch.empty(B, h, dtype=torch.bfloat16, device=device).data_ptr(), stage.logits.data_ptr(),
                        torch.empty(B, dtype=torch.int32, device=device).data_ptr(), ffi.current_stream(device)))
        torch.cuda.synchronize(device)
        self._launch = launch         # PipelineStage.capture_round re-captures the same launches into a round graph
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.device(device), torch.cuda.graph(self.graph):
            self._keep = launch()
        self.launches = ffi.lib.dli_last_launch_count() + (2 if self.streaming else 0)   # + eviction kernel, length reset
        self._len = gen.length        # host mirrors of the device-side length and position
        self._seen = gen.seen_tokens

    def before_replay(self) -> None:
        gen = self.cache.generation(self.gid)
        if self.streaming:
            if gen.length != self.cache.window_length or self.steps_left <= 0:
                raise RuntimeError("streaming graph: the window must be full and steps must remain "
                                   f"(length {gen.length}, steps left {self.steps_left})")
        elif gen.length + 1 > self.max_ctx:
            raise RuntimeError(f"graphed decode captured for at most {self.max_ctx} tokens")
        if gen.seen_tokens != self._seen:                              # eager steps ran in between: re-seed the position
            self.pos.fill_(gen.first_pos + gen.seen_tokens)

    def after_replay(self) -> None:
        gen = self.cache.generation(self.gid)
        if self.streaming:                                             # evicted one, appended one: the length stays W
            gen.evictions += 1
            self.steps_left -= 1
        else:
            gen.length += 1                                            # host mirror of the device-side add
        self._len = gen.length
        gen.seen_tokens += 1
        self._seen = gen.seen_tokens
        self.block.last_launch_count = self.launches

    def step(self, x: Optional[torch.Tensor] = None) -> torch.Tensor:
        self.before_replay()
        if x is not None and x.data_ptr() != self.x.data_ptr():
            self.x.copy_(x)
        self.graph.replay()
        self.after_replay()
        return self.out
