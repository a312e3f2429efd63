"""Minimal sequential client: embed -> spans -> final RMSNorm -> lm_head -> greedy argmax.

The reference has no client (model.py:19-23 keeps ``vocab_size``/``padding_idx`` but builds neither an
embedding nor a head; fact 7 of SURVEY.md) although its design — and BASELINE.json's north star —
assume one that "runs a sequential forward through spans".  This module is that client.  A span is
any callable with the reference's ``LlamaBlock.forward`` signature (model.py:25-34); a session is a
``generation_id``.  With bf16 CUDA tensors the client's tail -- final RMSNorm, vocabulary projection,
greedy argmax -- is one C call (``dli_lm_head_argmax``: RMSNorm kernel, the tcgen05 weight-streaming
GEMM over the ``[vocab, h]`` head, argmax kernel); in a pipeline the same three kernels close the
last stage's step (``parallel/pipeline.py``, token ring).  With CPU tensors (how
``tests/golden/make_golden.py`` drives the REFERENCE's spans in fp32 to produce the fixtures) the
tail is plain torch arithmetic -- that path exists for the fixture generator, not for serving.
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
        last = h[:, -1, :]
        if last.is_cuda:
            return self._head_cuda(last.contiguous())
        hn = _rmsnorm(last, self.norm, self.norm_eps)
        return hn @ self.lm_head.float().t()

    def _head_cuda(self, last: torch.Tensor) -> torch.Tensor:
        """fp32 logits ``[B, V]`` through ``dli_lm_head_argmax``; the kernel's greedy token ids are kept in
        ``self.last_tokens`` (int32 ``[B]``)."""
        from distributed_llm_inference_b200 import _ffi

        B, h = last.shape
        V = self.lm_head.shape[0]
        if last.dtype != torch.bfloat16 or self.lm_head.dtype != torch.bfloat16 or self.norm.dtype != torch.bfloat16:
            raise TypeError("the CUDA client tail needs bf16 hidden states, norm and lm_head")
        dev = last.device
        logits = torch.empty(B, V, dtype=torch.float32, device=dev)
        self.last_tokens = torch.empty(B, dtype=torch.int32, device=dev)
        xn = torch.empty(B, h, dtype=torch.bfloat16, device=dev)
        with torch.cuda.device(dev):
            _ffi.check(_ffi.lib.dli_lm_head_argmax(last.data_ptr(), 0, self.norm.data_ptr(), self.norm_eps,
                                                   self.lm_head.data_ptr(), V, h, B, xn.data_ptr(), logits.data_ptr(),
                                                   self.last_tokens.data_ptr(), _ffi.current_stream(dev)))
        return logits

    @torch.no_grad()
    def greedy(self, generation_id: str, prompt: torch.Tensor, n_new: int,
               forced: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, List[torch.Tensor]]:
        """Greedy decoding of ``n_new`` tokens.  ``forced [B, n_new]`` teacher-forces the token that
        is fed back (the argmax is still what gets recorded)."""
        logits = self.step(generation_id, prompt, 0)
        toks, logs = [], []
        pos = prompt.shape[1]
        for i in range(n_new):
            nxt = self.last_tokens.long() if logits.is_cuda else logits.argmax(-1)
            toks.append(nxt)
            logs.append(logits)
            if i == n_new - 1:
                break
            feed = forced[:, i].to(nxt.device) if forced is not None else nxt
            logits = self.step(generation_id, feed[:, None], pos)
            pos += 1
        return torch.stack(toks, dim=1), logs
