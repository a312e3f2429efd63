"""Paged, per-generation KV cache with the reference's ``PartialLlamaSinkCache`` interface.

Reference: distributed_llm_inference/models/llama/cache.py:7-135.  There, every generation owns Python
lists of ``[B, KV, S, hd]`` tensors that are re-allocated with ``torch.cat`` on every step (cache.py:107-109;
22 % of the reference's CPU decode time).  Here every layer owns a pool of fixed-size pages
``[n_pages, KV, page_tokens, hd]`` (bf16, keys stored rotated exactly as the reference stores them,
modules.py:76-85); a generation is a block table (one row of page ids per sequence) plus a length, both
mirrored on the device so that the kernels — and a replayed CUDA graph — read them from HBM.

Same constructor signature ``(window_length, num_sink_tokens)`` (cache.py:11), same errors
(``ValueError`` when the generation id is missing: cache.py:24-30,51-52,71-72), same
``get_seq_length(layer_idx, generation_id)`` (cache.py:50-62), many concurrent generations keyed by
``generation_id`` (cache.py:14-19).  Deviations, all deliberate:
  * the constructor works (the reference's raises, cache.py:12 / SURVEY D1);
  * layers are indexed by their LOCAL position in the span;
  * attention-sink eviction (cache.py:111-133) runs as one in-place kernel over all layers of the span
    (``evict`` -> ``dli_kv_evict``), for single tokens and for multi-token chunks alike (the chunk must leave room for
    at least one kept token: ``n < window - sinks``);
  * ``update()`` (the per-layer tensor API of cache.py:64-135) is provided for callers that hold K/V
    tensors themselves, eviction and cos/sin history included; the span forward does not use it — its fused
    kernels append in place.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Tuple

import torch


class _Generation:
    __slots__ = ("batch", "length", "pages", "block_table", "seq_lens", "max_pages", "seen_tokens", "first_pos",
                 "evictions", "cos_hist", "sin_hist")

    def __init__(self, batch: int, max_pages: int, device):
        self.batch = batch
        self.length = 0                      # tokens stored (same for every sequence of the generation)
        self.seen_tokens = 0                 # cache.py:86-87
        self.first_pos = 0                   # absolute position of the generation's first token (cos/sin history)
        self.evictions = 0
        self.cos_hist = None                 # update(): cos/sin rows of the generation's first `window` positions
        self.sin_hist = None                 # (cache.py:91-101); LlamaBlock.forward takes them from its rotary table
        self.pages: List[List[int]] = [[] for _ in range(batch)]
        self.max_pages = max_pages
        self.block_table = torch.zeros(batch, max_pages, dtype=torch.int32, device=device)
        self.seq_lens = torch.zeros(batch, dtype=torch.int32, device=device)


class PartialLlamaSinkCache:
    """Distributed sink cache: many generations at once, paged storage (reference cache.py:7-10)."""

    def __init__(self, window_length: int, num_sink_tokens: int, *, page_tokens: int = 16,
                 capacity_tokens: Optional[int] = None) -> None:
        if window_length <= 0 or num_sink_tokens < 0:
            raise ValueError("window_length must be positive and num_sink_tokens non-negative")
        self.window_length = window_length
        self.num_sink_tokens = num_sink_tokens
        self.page_tokens = page_tokens
        self.capacity_tokens = capacity_tokens
        self._gens: Dict[str, _Generation] = {}
        self._k: Optional[torch.Tensor] = None      # [n_layers, n_pages, KV, page_tokens, hd]
        self._v: Optional[torch.Tensor] = None
        self._free: List[int] = []
        self._shape: Optional[Tuple[int, int, int]] = None   # (n_layers, KV, hd)

    # ------------------------------------------------------------------ storage
    def bind(self, n_layers: int, n_kv_heads: int, head_dim: int, device, dtype=torch.bfloat16) -> None:
        """Allocate the page pools (once).  Called by ``LlamaBlock`` on first use; idempotent."""
        shape = (n_layers, n_kv_heads, head_dim)
        if self._k is not None:
            if self._shape != shape or self._k.device != torch.device(device):
                raise ValueError(f"cache already bound to {self._shape} on {self._k.device}")
            return
        cap = self.capacity_tokens if self.capacity_tokens is not None else min(self.window_length, 4096) * 4
        n_pages = max(1, -(-cap // self.page_tokens))
        # zero-filled: the tensor-core attention loads whole pages, and never-written slots must hold finite values
        # (their probabilities are exactly 0, but 0 * NaN would poison the P.V product)
        self._k = torch.zeros(n_layers, n_pages, n_kv_heads, self.page_tokens, head_dim, dtype=dtype, device=device)
        self._v = torch.zeros_like(self._k)
        self._free = list(range(n_pages - 1, -1, -1))
        self._shape = shape

    @property
    def n_pages(self) -> int:
        return 0 if self._k is None else self._k.shape[1]

    @property
    def free_pages(self) -> int:
        return len(self._free)

    def _max_pages_per_seq(self, min_tokens: int = 0) -> int:
        # the first chunk of a generation is stored whole even when it is longer than the window (cache.py:103-105)
        need = max(-(-self.window_length // self.page_tokens) + 1, -(-min_tokens // self.page_tokens))
        return min(need, max(1, self.n_pages))

    # ------------------------------------------------------------------ generations
    def open(self, generation_id: str, batch: int, min_tokens: int = 0) -> _Generation:
        if not generation_id:
            raise ValueError("generation_id not provided")
        if self._k is None:
            raise RuntimeError("cache is not bound to a span yet")
        g = self._gens.get(generation_id)
        if g is None:
            g = _Generation(batch, self._max_pages_per_seq(min_tokens), self._k.device)
            self._gens[generation_id] = g
        elif g.batch != batch:
            raise ValueError(f"generation {generation_id!r} was opened with batch {g.batch}, got {batch}")
        return g

    def close(self, generation_id: str) -> None:
        """Return the generation's pages to the pool (the reference never frees; sessions end with the process)."""
        g = self._gens.pop(generation_id, None)
        if g is not None:
            for row in g.pages:
                self._free.extend(row)

    def reserve(self, generation_id: str, batch: int, total_tokens: int) -> _Generation:
        """Make sure pages exist for ``total_tokens`` tokens per sequence (so that a captured graph can be
        replayed without touching the block table)."""
        g = self.open(generation_id, batch, total_tokens)
        if total_tokens > self.window_length and g.length > 0:
            raise ValueError(f"generation {generation_id!r}: {total_tokens} tokens exceed window_length="
                             f"{self.window_length}; evict first (PartialLlamaSinkCache.evict)")
        need = -(-total_tokens // self.page_tokens)
        if need > g.max_pages:
            raise ValueError(f"{total_tokens} tokens exceed the block table ({g.max_pages} pages of {self.page_tokens})")
        have = len(g.pages[0])
        if need > have:
            n_new = (need - have) * g.batch
            if n_new > len(self._free):
                raise RuntimeError(f"KV page pool exhausted: need {n_new} pages, {len(self._free)} free "
                                   f"(capacity_tokens={self.n_pages * self.page_tokens})")
            for b in range(g.batch):
                g.pages[b].extend(self._free.pop() for _ in range(need - have))
            host = torch.tensor([row + [0] * (g.max_pages - len(row)) for row in g.pages], dtype=torch.int32)
            g.block_table.copy_(host, non_blocking=False)
        return g

    def needs_eviction(self, generation_id: str, n_new: int) -> bool:
        """cache.py:107: append while ``n + len < window``, otherwise evict (a first chunk is always stored whole)."""
        g = self._gens.get(generation_id)
        return g is not None and g.length > 0 and n_new + g.length >= self.window_length

    def evict(self, generation_id: str, n_new: int, cos_hist: torch.Tensor, sin_hist: torch.Tensor) -> None:
        """Attention-sink eviction for all layers (cache.py:111-133): keep the first ``num_sink_tokens`` tokens and
        the last ``window - sinks - n_new``, re-rotating the kept keys back by ``n_new`` positions.  ``cos_hist`` /
        ``sin_hist``: fp32 ``[window, hd]`` rows of the generation's first ``window`` positions (the reference's
        ``_cos_cache`` / ``_sin_cache``).  Afterwards the generation holds ``window - n_new`` tokens."""
        import ctypes as C

        from distributed_llm_inference_b200 import _ffi

        g = self._gens[generation_id]
        W, s = self.window_length, self.num_sink_tokens
        if W - s - n_new <= 0:
            raise NotImplementedError(f"chunk of {n_new} tokens does not fit window {W} with {s} sink tokens")
        if cos_hist.shape[0] < W or not cos_hist.is_contiguous() or cos_hist.dtype != torch.float32:
            raise ValueError("cos/sin history must be contiguous fp32 [window, head_dim]")
        n_layers, KV, hd = self._shape
        kv = self.kv_desc(g, g.length)
        _ffi.check(_ffi.lib.dli_kv_evict(C.byref(kv), n_layers, g.batch, KV, hd, g.length, W, s, n_new,
                                         cos_hist.data_ptr(), sin_hist.data_ptr(), _ffi.current_stream()))
        g.length = W - n_new
        g.seq_lens.fill_(W - n_new)
        g.evictions += 1

    def advance(self, generation_id: str, n_tokens: int, local_first_layer: bool = True) -> None:
        g = self._gens[generation_id]
        g.length += n_tokens
        g.seq_lens.add_(n_tokens)
        if local_first_layer:
            g.seen_tokens += n_tokens

    def generation(self, generation_id: str) -> Optional[_Generation]:
        return self._gens.get(generation_id)

    # ------------------------------------------------------------------ reference API
    def get_seq_length(self, layer_idx: Optional[int] = 0, generation_id: Optional[str] = None) -> int:
        """cache.py:50-62."""
        if not generation_id:
            raise ValueError("generation_id not provided")
        g = self._gens.get(generation_id)
        if g is None:
            return 0
        if self._shape is not None and layer_idx is not None and layer_idx >= self._shape[0]:
            return 0
        return g.length

    def get_max_length(self) -> Optional[int]:
        return self.window_length

    def update(self, key_states: torch.Tensor, value_states: torch.Tensor, layer_idx: int,
               cache_kwargs: Optional[Dict[str, Any]] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """Per-layer tensor API of cache.py:64-135 for callers that hold rotated K and V ``[B, KV, n, hd]``
        themselves: appends them to the pages of local layer ``layer_idx`` and returns the gathered
        ``[B, KV, S, hd]`` views' copies.  The last layer of the span advances the generation length."""
        if cache_kwargs is None or "generation_id" not in cache_kwargs:
            raise ValueError("generation_id not found in cache_kwargs")
        gid = cache_kwargs.get("generation_id")
        B, KV, n, hd = key_states.shape
        if self._k is None:
            raise RuntimeError("cache is not bound to a span yet (call bind(n_layers, n_kv_heads, head_dim, device))")
        sin, cos = cache_kwargs.get("sin"), cache_kwargs.get("cos")
        n_layers = self._shape[0]
        g = self._gens.get(gid)
        if layer_idx == 0:
            if cos is not None and sin is not None:          # cache.py:91-101: history until `window` rows
                g0 = self.open(gid, B, n)
                c2, s2 = (cos, sin) if cos.dim() == 2 else (cos[0], sin[0])
                c2, s2 = c2.to(self._k.device, torch.float32), s2.to(self._k.device, torch.float32)
                if cos.dim() == 2 or g0.cos_hist is None:
                    g0.cos_hist, g0.sin_hist = c2.contiguous(), s2.contiguous()
                elif g0.cos_hist.shape[0] < self.window_length:
                    g0.cos_hist = torch.cat([g0.cos_hist, c2], dim=0)
                    g0.sin_hist = torch.cat([g0.sin_hist, s2], dim=0)
            if self.needs_eviction(gid, n):                  # cache.py:111-133, all layers of the span in one launch
                g = self._gens[gid]
                W = self.window_length
                if g.cos_hist is None or g.cos_hist.shape[0] < W:
                    raise ValueError("eviction needs the cos/sin history of the first `window_length` positions: pass "
                                     "cos and sin in cache_kwargs on every update of layer 0 (cache.py:91-101)")
                self.evict(gid, n, g.cos_hist[:W].contiguous(), g.sin_hist[:W].contiguous())
        g = self.reserve(gid, B, (self._gens[gid].length if gid in self._gens else 0) + n)
        pt = self.page_tokens
        pos = torch.arange(g.length, g.length + n, device=self._k.device)
        table = g.block_table.long()
        page = table[:, pos // pt]                              # [B, n]
        slot = (pos % pt)[None, :].expand(B, n)
        self._k[layer_idx][page, :, slot] = key_states.transpose(1, 2).to(self._k.dtype)
        self._v[layer_idx][page, :, slot] = value_states.transpose(1, 2).to(self._v.dtype)
        total = g.length + n
        allpos = torch.arange(total, device=self._k.device)
        pg, sl = table[:, allpos // pt], (allpos % pt)[None, :].expand(B, total)
        k_all = self._k[layer_idx][pg, :, sl].transpose(1, 2)
        v_all = self._v[layer_idx][pg, :, sl].transpose(1, 2)
        if layer_idx == n_layers - 1:
            self.advance(gid, n)
        return k_all, v_all

    # ------------------------------------------------------------------ kernel-side view
    def kv_desc(self, g: _Generation, max_ctx: int):
        from distributed_llm_inference_b200 import _ffi

        k = self._k
        return _ffi.KvDesc(k.data_ptr(), self._v.data_ptr(), k.stride(0) * k.element_size(), self.page_tokens,
                           k.shape[1], g.block_table.data_ptr(), g.seq_lens.data_ptr(), g.max_pages, max_ctx)
