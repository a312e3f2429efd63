"""Span-to-span hand-off on one NVSwitch box: one pipeline stage per GPU, NCCL P2P instead of a network hop.

In the reference, consecutive spans live on different servers and the hidden states travel through hivemind
(serialize -> socket -> deserialize; only imported at server/backend.py:4-7, never exercised by code in the tree).
Here stage ``r`` owns layers ``[r*n, (r+1)*n)`` and their KV pages; the ``[B, L, h]`` bf16 hand-off tensor goes to
stage ``r+1`` with ``torch.distributed`` ``isend/irecv`` (NCCL over NVLink 5; gloo on CPU in the tests), and the
last stage closes the ring back to stage 0 (where the client would sample and embed the next token).  Several
micro-batches are interleaved so that every stage is busy once the pipe is full: steady-state utilisation is
``min(1, micro_batches / stages)``.

The exchange is exactly one send/recv per stage boundary per micro-batch; there is no all-reduce / all-gather
(the reference has no tensor parallelism).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


class PipelineStage:
    def __init__(self, span, cache, rank: int, world: int, device, hidden_size: Optional[int] = None,
                 dtype=torch.bfloat16):
        self.span, self.cache = span, cache
        self.rank, self.world, self.device = rank, world, device
        self.hidden_size = hidden_size if hidden_size is not None else span.config.hidden_size
        self.dtype = dtype
        self.launches_per_round = 0
        self._pending = []          # (work, tensor) of sends still in flight: keep tensors alive

    # -- P2P ------------------------------------------------------------------------------------------
    def _send(self, t: torch.Tensor, dst: int) -> None:
        self._pending.append((dist.isend(t, dst), t))

    def _recv(self, shape, src: int) -> torch.Tensor:
        buf = torch.empty(shape, dtype=self.dtype, device=self.device)
        dist.irecv(buf, src).wait()       # NCCL: the current stream waits for the transfer, the host does not
        return buf

    def _reap(self) -> None:
        for work, _ in self._pending:
            work.wait()
        self._pending.clear()

    # -- one round: every micro-batch passes through this stage once --------------------------------------
    def run_round(self, generation_ids: Sequence[str], inputs: Optional[List[torch.Tensor]], batch: int, q_len: int,
                  start_pos: int, feed_back: bool = True) -> Optional[List[torch.Tensor]]:
        """Stage 0 consumes ``inputs`` (one ``[B, q_len, h]`` tensor per micro-batch); other stages receive theirs
        from the previous stage.  Returns, on stage 0, the tensors that came back from the last stage (the next
        round's inputs when ``feed_back``); ``None`` elsewhere."""
        first, last = self.rank == 0, self.rank == self.world - 1
        self._reap()
        cp = torch.arange(start_pos, start_pos + q_len, device=self.device)
        shape = (batch, q_len, self.hidden_size)
        outs, launches = [], 0
        for m, gid in enumerate(generation_ids):
            x = inputs[m] if first else self._recv(shape, self.rank - 1)
            y = self.span(gid, x, past_key_value=self.cache, cache_position=cp)[0]
            launches += getattr(self.span, "last_launch_count", 0)
            if not last:
                self._send(y, self.rank + 1)
            elif self.world > 1 and feed_back:
                self._send(y, 0)
            else:
                outs.append(y)
        self.launches_per_round = launches
        if self.world == 1:
            return outs
        if first and feed_back:
            return [self._recv(shape, self.world - 1) for _ in generation_ids]
        return outs if last else None
