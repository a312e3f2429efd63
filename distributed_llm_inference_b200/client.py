"""Minimal sequential client: embed -> spans -> final RMSNorm -> lm_head -> greedy argmax.

The reference has no client (model.py:19-23 keeps ``vocab_size``/``padding_idx`` but builds neither an
embedding nor a head; fact 7 of SURVEY.md) although its design — and BASELINE.json's north star —
assume one that "runs a sequential forward through spans".  This module is that client.  A span is
any callable with the reference's ``LlamaBlock.forward`` signature (model.py:25-34); a session is a
``generation_id``.  The embedding lookup, final norm and vocabulary projection run as plain torch
ops on whatever device the tensors live on: they are the caller side of the hot path (SURVEY
section 8f, rank 1), not part of the span.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F


def _rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    x32 = x.float()
    y = x32 * torch.rsqrt(x32.pow(2).mean(-1, keepdim=True) + eps)
    return (w.float() * y)


class SequentialClient:
    def __init__(self, embed_tokens: torch.Tensor, norm: torch.Tensor, lm_head: torch.Tensor,
                 spans: Sequence, caches: Sequence, norm_eps: float = 1e-6,
                 hand_off=None):
        """``hand_off(h, i)`` (optional) is applied to the hidden states between span ``i`` and ``i+1``
        (e.g. a device-to-device copy); identity by default."""
        self.embed_tokens, self.norm, self.lm_head = embed_tokens, norm, lm_head
        self.spans, self.caches = list(spans), list(caches)
        self.norm_eps = norm_eps
        self.hand_off = hand_off

    @torch.no_grad()
    def step(self, generation_id: str, token_ids: torch.Tensor, start_pos: int) -> torch.Tensor:
        """Feed ``token_ids [B, L]`` at absolute positions ``start_pos..``; return fp32 logits ``[B, V]``
        of the last position."""
        h = F.embedding(token_ids, self.embed_tokens)
        cp = torch.arange(start_pos, start_pos + token_ids.shape[1], device=h.device)
        for i, (span, cache) in enumerate(zip(self.spans, self.caches)):
            h = span(generation_id, h, past_key_value=cache, cache_position=cp)[0]
            if self.hand_off is not None and i + 1 < len(self.spans):
                h = self.hand_off(h, i)
        hn = _rmsnorm(h[:, -1, :], self.norm, self.norm_eps)
        return hn @ self.lm_head.float().t()

    @torch.no_grad()
    def greedy(self, generation_id: str, prompt: torch.Tensor, n_new: int,
               forced: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, List[torch.Tensor]]:
        """Greedy decoding of ``n_new`` tokens.  ``forced [B, n_new]`` teacher-forces the token that
        is fed back (the argmax is still what gets recorded)."""
        logits = self.step(generation_id, prompt, 0)
        toks, logs = [], []
        pos = prompt.shape[1]
        for i in range(n_new):
            nxt = logits.argmax(-1)
            toks.append(nxt)
            logs.append(logits)
            if i == n_new - 1:
                break
            feed = forced[:, i].to(nxt.device) if forced is not None else nxt
            logits = self.step(generation_id, feed[:, None], pos)
            pos += 1
        return torch.stack(toks, dim=1), logs
