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
        self.graphs = {}            # generation id -> GraphedDecode (whole stage step as one CUDA graph)
        self.handoff = "nccl" if world > 1 else "local"
        self.handoff_error = None
        self._ring_pending = False
        self.trace = None           # set to [] to collect (recv_wait_ms, step_ms) CUDA-event pairs per micro-batch
        self._round = None          # (generation ids, CUDA graph of a whole round, kept tensors): capture_round()
        self.round_error = None

    def enable_graphs(self, generation_ids: Sequence[str], max_ctx: int, batch: Optional[int] = None,
                      peer_store: bool = True) -> str:
        """Capture this stage's decode step of every (prefilled) micro-batch into a CUDA graph.

        ``peer_store`` (multi-GPU, NVLink/NVSwitch): every stage owns an "inbox" ``[micro_batches, B, 1, h]`` in
        symmetric memory (peer-mapped by all ranks).  Stage r's graph for micro-batch m reads its input from its own
        inbox slot m and its LAST KERNEL writes the hand-off tensor directly into slot m of stage r+1's inbox over
        NVLink — no send/recv kernel, no extra copy, nothing competing for SMs on a second stream; a one-word
        semaphore per (edge, micro-batch) orders producer and consumer.  If symmetric memory cannot be set up the
        stage falls back to NCCL ``isend/irecv`` into the graphs' static buffers.  Returns the hand-off mode."""
        self.handoff = "local"
        peer = None
        if self.world > 1 and peer_store:
            try:
                import torch.distributed._symmetric_memory as symm_mem

                if batch is None:
                    batch = self.cache.generation(generation_ids[0]).batch
                shape = (len(generation_ids), batch, 1, self.hidden_size)
                if hasattr(symm_mem, "enable_symm_mem_for_group"):
                    try:
                        symm_mem.enable_symm_mem_for_group(dist.group.WORLD.group_name)
                    except Exception:
                        pass
                self.inbox = symm_mem.empty(shape, dtype=self.dtype, device=self.device)
                self.inbox.zero_()
                self.hdl = symm_mem.rendezvous(self.inbox, dist.group.WORLD)
                peer = self.hdl.get_buffer((self.rank + 1) % self.world, shape, self.dtype)
                self.handoff = "peer_store"
            except Exception as e:  # noqa: BLE001 - any failure means: use NCCL
                self.handoff_error = repr(e)
                peer = None
        if self.world > 1 and peer is None:
            self.handoff = "nccl"
        ok = torch.tensor([1 if self.handoff != "nccl" else 0], device=self.device)
        if self.world > 1:          # all stages must agree on the protocol
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if ok.item() == 0:
                self.handoff, peer = "nccl", None
        for m, gid in enumerate(generation_ids):
            if peer is not None:
                self.graphs[gid] = self.span.graphed_decode(gid, self.cache, max_ctx, x=self.inbox[m], out=peer[m])
            else:
                self.graphs[gid] = self.span.graphed_decode(gid, self.cache, max_ctx)
        if self.world > 1:
            dist.barrier()
        return self.handoff

    def capture_round(self, generation_ids: Sequence[str]) -> bool:
        """Peer-store hand-off only: capture a whole steady-state round -- for every micro-batch ``wait(prev) -> stage
        step -> signal(next)`` -- into ONE CUDA graph, so a round costs one graph launch instead of one per micro-batch
        plus two semaphore launches each.  All stages must call this together; if any of them cannot capture, none uses
        it.  The per-micro-batch graphs stay in place for the first round (fresh inputs) and for tracing."""
        ok = 1 if (self.handoff == "peer_store" and all(g in self.graphs for g in generation_ids)) else 0
        graph, keep = None, []
        if ok:
            try:
                prev, nxt = (self.rank - 1) % self.world, (self.rank + 1) % self.world
                torch.cuda.synchronize(self.device)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    for m, gid in enumerate(generation_ids):
                        self.hdl.wait_signal(prev, m, 20000)
                        keep.append(self.graphs[gid]._launch())
                        self.hdl.put_signal(nxt, m, 20000)
            except Exception as e:  # noqa: BLE001 - any failure means: keep the per-micro-batch graphs
                self.round_error = repr(e)
                ok = 0
        flag = torch.tensor([ok], device=self.device)
        if self.world > 1:
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        self._round = (tuple(generation_ids), graph, keep) if flag.item() == 1 else None
        return self._round is not None

    # -- P2P ------------------------------------------------------------------------------------------
    def _send(self, t: torch.Tensor, dst: int) -> None:
        self._pending.append((dist.isend(t, dst), t))

    def _recv(self, shape, src: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        buf = out if out is not None else torch.empty(shape, dtype=self.dtype, device=self.device)
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
        if q_len == 1 and self.graphs:
            return self._run_round_graphed(generation_ids, inputs, batch, feed_back)
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

    def _run_round_graphed(self, generation_ids, inputs, batch, feed_back):
        if self.handoff == "peer_store":
            return self._run_round_peer_store(generation_ids, inputs, feed_back)
        first, last = self.rank == 0, self.rank == self.world - 1
        shape = (batch, 1, self.hidden_size)
        outs, launches = [], 0
        for m, gid in enumerate(generation_ids):
            g = self.graphs[gid]
            if self.trace is not None:
                ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                ev[0].record()
            if first:
                if inputs[m].data_ptr() != g.x.data_ptr():
                    g.x.copy_(inputs[m])
            else:
                self._recv(shape, self.rank - 1, out=g.x)
            if self.trace is not None:
                ev[1].record()
            y = g.step()
            if self.trace is not None:
                ev[2].record()
                self.trace.append(ev)
            launches += g.launches
            if not last:
                self._send(y, self.rank + 1)
            elif self.world > 1 and feed_back:
                self._send(y, 0)
            else:
                outs.append(y)
        self.launches_per_round = launches
        if self.world == 1:
            return outs
        if first and feed_back:   # ring closure straight into the graphs' input buffers
            return [self._recv(shape, self.world - 1, out=self.graphs[gid].x) for gid in generation_ids]
        return outs if last else None

    def _run_round_peer_store(self, generation_ids, inputs, feed_back):
        """wait(prev, m) -> graph m (last kernel stores into the next stage's inbox) -> signal(next, m), all on one
        stream.  A slot is never overwritten early: the producer's next write to it depends, through the ring, on the
        consumer's previous step."""
        first, last = self.rank == 0, self.rank == self.world - 1
        prev, nxt = (self.rank - 1) % self.world, (self.rank + 1) % self.world
        launches = 0
        if (self._round is not None and feed_back and self.trace is None and tuple(generation_ids) == self._round[0]
                and (not first or self._ring_pending)):
            gs = [self.graphs[gid] for gid in generation_ids]
            for g in gs:
                g.before_replay()
            self._round[1].replay()
            for g in gs:
                g.after_replay()
            self.launches_per_round = sum(g.launches + 2 for g in gs)
            return [g.x for g in gs] if first else None
        for m, gid in enumerate(generation_ids):
            g = self.graphs[gid]
            if self.trace is not None:
                ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                ev[0].record()
            if first and not self._ring_pending:
                if inputs[m].data_ptr() != g.x.data_ptr():
                    g.x.copy_(inputs[m])                   # fresh input injected at the head of the pipe
            else:
                self.hdl.wait_signal(prev, m, 20000)
            if self.trace is not None:
                ev[1].record()
            g.step()
            if not last or feed_back:
                self.hdl.put_signal(nxt, m, 20000)
            if self.trace is not None:
                ev[2].record()
                self.trace.append(ev)
            launches += g.launches + 2
        self.launches_per_round = launches
        if first:
            self._ring_pending = feed_back
        if first and feed_back:
            return [self.graphs[gid].x for gid in generation_ids]   # filled by the last stage; waited on next round
        return None

    def drain(self, n_micro_batches: int) -> None:
        """Stage 0: consume the ring-closure signals of the last round (before switching hand-off mode)."""
        if self.handoff == "peer_store" and self.rank == 0 and self._ring_pending:
            for m in range(n_micro_batches):
                self.hdl.wait_signal(self.world - 1, m, 20000)
            self._ring_pending = False
