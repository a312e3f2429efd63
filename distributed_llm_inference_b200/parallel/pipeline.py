"""Span-to-span hand-off on one NVSwitch box: one pipeline stage per GPU instead of a network hop.

In the reference, consecutive spans live on different servers and the hidden states travel through hivemind
(serialize -> socket -> deserialize; only imported at server/backend.py:4-7, never exercised by code in the tree).
Here stage ``r`` owns layers ``[r*n, (r+1)*n)`` and their KV pages, and several micro-batches are interleaved so that
every stage is busy once the pipe is full: steady-state utilisation is ``min(1, micro_batches / stages)``.  The exchange
is exactly one hand-off per stage boundary per micro-batch; there is no all-reduce / all-gather (the reference has no
tensor parallelism).

Decode hand-off (``enable_graphs``): every stage owns, in symmetric (peer-mapped) memory, an inbox slot and a flag per
micro-batch.  A stage step is ONE C call (``dli_stage_decode_step``) whose own kernels do the hand-off: the entry kernel
spins on the stage's flag until the producer has stored the step's input, the exit kernel stores the ``[B, 1, h]``
tensor straight into the next stage's inbox over NVLink, fences, and releases that stage's flag -- no send/recv kernel, no
separate semaphore launches.  Two kinds of ring:

``hidden``  the last stage stores its output hidden states into stage 0's inbox (what the benchmark's device-timed
            ``value`` measures: the spans alone, as the reference defines them -- it has no embedding and no head);
``tokens``  SURVEY 8e: stage 0 looks the step's input up in the embedding table by device-resident token ids, the
            last stage ends in final RMSNorm -> lm_head -> greedy argmax and stores token ids ``[B]`` into stage 0's
            token inbox: the sequential client of the north star, closed on the GPUs.

A whole steady-state round (every micro-batch once) is replayed as one CUDA graph per stage.  Prefill chunks and the
fallback when symmetric memory is unavailable use ``torch.distributed`` ``isend/irecv`` (NCCL; gloo on CPU in tests).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

# Tokens per chunk when ONE long prompt is prefilled through the stages (run_round with q_len = chunk, start_pos advancing).
# N identical stages and chunk times t_i take sum(t_i) + (N - 1) * mean(t_i): a stage's own time for an 8k prompt is flat
# from 512 to 2048 tokens per chunk and rises below 512 (the 256 x 256 CTA-pair GEMM tiles stop filling the SMs), so 512
# minimises the sum (70B, 8192 tokens: 211 ms on 8 stages against 396 ms with 2048-token chunks).
PREFILL_CHUNK_TOKENS = 512


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


class _Ring:
    """Buffers, flags and captured step graphs of one ring kind on one stage."""

    def __init__(self, kind: str):
        self.kind = kind
        self.graphs: Dict[str, object] = {}     # generation id -> GraphedDecode
        self.round = None                        # (generation ids, CUDAGraph, kept objects)
        self.steps = 0                           # stage steps launched per micro-batch so far (host mirror of state[:, 0])
        self.gids: Sequence[str] = ()


class PipelineStage:
    def __init__(self, span, cache, rank: int, world: int, device, hidden_size: Optional[int] = None,
                 dtype=torch.bfloat16, client: Optional[dict] = None):
        """``client`` (token ring only): ``{"embed_tokens": [V, h]}`` on stage 0 and ``{"norm": [h], "lm_head": [V, h],
        "norm_eps": float}`` on the last stage, bf16 on this stage's device."""
        self.span, self.cache = span, cache
        self.rank, self.world, self.device = rank, world, device
        self.hidden_size = hidden_size if hidden_size is not None else span.config.hidden_size
        self.dtype = dtype
        self.client = client or {}
        self.launches_per_round = 0
        self._pending = []          # (work, tensor) of sends still in flight: keep tensors alive
        self.rings: Dict[str, _Ring] = {}
        self.handoff = "nccl" if world > 1 else "local"
        self.handoff_error = None
        self.round_error = None
        # chain the steps of different micro-batches with programmatic dependent launch (the next step's entry kernel and
        # its first GEMM's weight ring are resident while the current step drains) instead of a stream barrier in front of
        # every step: 8 stages x 4 layers of the 7B shape, 3.311 -> 3.282 ms per round; no difference at 2 stages
        self.chain_steps = True
        self.trace = None           # set to [] to collect (start, end) CUDA-event pairs per micro-batch (per-step graphs)

    # back-compat views used by bench.py / tests: the hidden ring's graphs
    @property
    def graphs(self):
        r = self.rings.get("hidden")
        return r.graphs if r is not None else {}

    # ------------------------------------------------------------------------------------------ set-up
    def _alloc_ring(self, ring: _Ring, n_mb: int, batch: int, peer_store: bool) -> bool:
        """Inbox, token inbox and flags: symmetric memory when there are peers, plain device memory on one GPU."""
        h = self.hidden_size
        off_flags = 0
        off_tok = _align(n_mb * 64)                                   # one 64-byte line per flag
        off_inbox = off_tok + _align(n_mb * batch * 4)
        total = off_inbox + _align(n_mb * batch * h * 2)
        peers = {}
        if self.world > 1:
            if not peer_store:
                return False
            try:
                import torch.distributed._symmetric_memory as symm_mem

                if hasattr(symm_mem, "enable_symm_mem_for_group"):
                    try:
                        symm_mem.enable_symm_mem_for_group(dist.group.WORLD.group_name)
                    except Exception:
                        pass
                buf = symm_mem.empty((total,), dtype=torch.uint8, device=self.device)
                buf.zero_()
                hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
                for p in {(self.rank + 1) % self.world, 0}:
                    peers[p] = hdl.get_buffer(p, (total,), torch.uint8)
                ring.hdl = hdl
            except Exception as e:  # noqa: BLE001 - any failure means: use NCCL
                self.handoff_error = repr(e)
                return False
        else:
            buf = torch.zeros(total, dtype=torch.uint8, device=self.device)
            peers[0] = buf

        def carve(b):
            return (b[off_flags:off_flags + n_mb * 64].view(torch.int32).view(n_mb, 16),
                    b[off_tok:off_tok + n_mb * batch * 4].view(torch.int32).view(n_mb, batch),
                    b[off_inbox:off_inbox + n_mb * batch * h * 2].view(torch.bfloat16).view(n_mb, batch, 1, h))

        ring.buf = buf
        ring.flags, ring.tok_inbox, ring.inbox = carve(buf)
        ring.peer = {p: carve(b) for p, b in peers.items()}
        ring.state = torch.zeros(n_mb, 4, dtype=torch.int32, device=self.device)
        return True

    def enable_graphs(self, generation_ids: Sequence[str], max_ctx: int, batch: Optional[int] = None,
                      peer_store: bool = True, ring: str = "hidden") -> str:
        """Capture this stage's decode step of every (prefilled) micro-batch into a CUDA graph and set up the hand-off.
        All stages call this together.  Returns the hand-off mode: ``peer_store`` (in-kernel flags over peer memory),
        ``local`` (one GPU) or ``nccl`` (fallback: graphs with static buffers, isend/irecv between them)."""
        from distributed_llm_inference_b200.models.llama.model import StageIO

        if ring not in ("hidden", "tokens"):
            raise ValueError(ring)
        r = _Ring(ring)
        first, last = self.rank == 0, self.rank == self.world - 1
        if batch is None:
            batch = self.cache.generation(generation_ids[0]).batch
        n_mb = len(generation_ids)
        ok = self._alloc_ring(r, n_mb, batch, peer_store)
        flag = torch.tensor([1 if ok else 0], device=self.device)
        if self.world > 1:          # all stages must agree on the protocol
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            ok = flag.item() == 1
        if ring == "tokens" and not ok:
            raise RuntimeError(f"the token ring needs peer-mapped memory: {self.handoff_error}")
        r.gids = tuple(generation_ids)
        if not ok:
            self.handoff = "nccl"
            for gid in generation_ids:
                r.graphs[gid] = self.span.graphed_decode(gid, self.cache, max_ctx)
        else:
            self.handoff = "peer_store" if self.world > 1 else "local"
            nxt = (self.rank + 1) % self.world
            tokens = ring == "tokens"
            if tokens and last:
                V = self.client["lm_head"].shape[0]
                r.logits = torch.zeros(batch, V, dtype=torch.float32, device=self.device)
            for m, gid in enumerate(generation_ids):
                p_flags, p_tok, p_inbox = r.peer[nxt]
                io = StageIO(
                    hidden_in=None if (tokens and first) else r.inbox[m],
                    token_ids=r.tok_inbox[m] if (tokens and first) else None,
                    embed_table=self.client["embed_tokens"] if (tokens and first) else None,
                    hidden_out=None if (tokens and last) else p_inbox[m],
                    final_norm=self.client["norm"] if (tokens and last) else None,
                    lm_head=self.client["lm_head"] if (tokens and last) else None,
                    logits=r.logits if (tokens and last) else None,
                    tokens_out=p_tok[m] if (tokens and last) else None,
                    eps_final=float(self.client.get("norm_eps", 1e-6)),
                    wait_flag=r.flags[m], signal_flag=p_flags[m], state=r.state[m],
                    wait_lag=1 if first else 0,
                    chain=1 if (n_mb > 1 and self.chain_steps) else 0)       # consecutive steps in the stream belong to different generations
                r.graphs[gid] = self.span.graphed_decode(gid, self.cache, max_ctx, stage=io)
        self.rings[ring] = r
        if self.world > 1:
            torch.cuda.synchronize(self.device)
            dist.barrier()
        return self.handoff

    def capture_round(self, generation_ids: Sequence[str], ring: str = "hidden") -> bool:
        """Peer-store hand-off only: capture a whole steady-state round -- every micro-batch's stage step, chained --
        into ONE CUDA graph, so a round costs one graph launch.  All stages must call this together; if any of them
        cannot capture, none uses it.  The per-micro-batch graphs stay in place (tracing, host-fed rounds)."""
        r = self.rings.get(ring)
        ok = 1 if (r is not None and self.handoff in ("peer_store", "local")
                   and all(g in r.graphs for g in generation_ids)) else 0
        graph, keep = None, []
        if ok:
            try:
                torch.cuda.synchronize(self.device)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    for gid in generation_ids:
                        keep.append(r.graphs[gid]._launch())
            except Exception as e:  # noqa: BLE001 - any failure means: keep the per-micro-batch graphs
                self.round_error = repr(e)
                ok = 0
        flag = torch.tensor([ok], device=self.device)
        if self.world > 1:
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if r is not None:
            r.round = (tuple(generation_ids), graph, keep) if flag.item() == 1 else None
        return r is not None and r.round is not None

    @property
    def _round(self):               # back-compat for tests
        r = self.rings.get("hidden")
        return None if r is None else r.round

    # -- P2P (prefill chunks, NCCL fallback) --------------------------------------------------------------
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
                  start_pos: int, feed_back: bool = True, ring: str = "hidden") -> Optional[List[torch.Tensor]]:
        """Stage 0 consumes ``inputs`` (one tensor per micro-batch: ``[B, q_len, h]`` hidden states, or ``[B]`` int32
        token ids for the token ring); other stages receive theirs from the previous stage.  Returns, on stage 0, what
        came back from the last stage (the next round's inputs when ``feed_back``); ``None`` elsewhere.  Graphed decode
        rounds over peer memory always close the ring; pass ``inputs=None`` on stage 0 to let them feed themselves."""
        first, last = self.rank == 0, self.rank == self.world - 1
        self._reap()
        r = self.rings.get(ring)
        if q_len == 1 and r is not None and r.graphs:
            if self.handoff == "nccl":
                return self._run_round_graphed_nccl(r, generation_ids, inputs, batch, feed_back)
            return self._run_round_flags(r, generation_ids, inputs)
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

    def prefill(self, generation_ids: Sequence[str], prompts: Optional[List[torch.Tensor]], batch: int, prompt_len: int,
                chunk_tokens: Optional[int] = None) -> Optional[List[torch.Tensor]]:
        """Prompts of ``prompt_len`` tokens through the stages in chunks of ``chunk_tokens`` (default
        ``PREFILL_CHUNK_TOKENS``; the whole prompt when it is shorter): stage s works on chunk c while stage s + 1 works
        on chunk c - 1.  ``prompts`` (stage 0 only): one ``[B, prompt_len, h]`` tensor per micro-batch.  The last stage
        returns the hidden states of the LAST chunk per micro-batch (what the head needs), the others ``None``."""
        chunk = max(1, min(chunk_tokens or PREFILL_CHUNK_TOKENS, prompt_len))
        out = None
        for c0 in range(0, prompt_len, chunk):
            n = min(chunk, prompt_len - c0)
            ins = [p[:, c0:c0 + n].contiguous() for p in prompts] if self.rank == 0 else None
            out = self.run_round(generation_ids, ins, batch, n, c0, feed_back=False)
        return out

    def _run_round_graphed_nccl(self, r, generation_ids, inputs, batch, feed_back):
        first, last = self.rank == 0, self.rank == self.world - 1
        shape = (batch, 1, self.hidden_size)
        outs, launches = [], 0
        for m, gid in enumerate(generation_ids):
            g = r.graphs[gid]
            if first:
                if inputs[m].data_ptr() != g.x.data_ptr():
                    g.x.copy_(inputs[m])
            else:
                self._recv(shape, self.rank - 1, out=g.x)
            y = g.step()
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
            return [self._recv(shape, self.world - 1, out=r.graphs[gid].x) for gid in generation_ids]
        return outs if last else None

    def inject(self, ring: str, m: int, x: torch.Tensor) -> None:
        """Stage 0: put a fresh input (hidden states or token ids) at the head of the pipe for micro-batch ``m``.  The
        slot is also the target of the ring closure, so the copy is ordered behind the arrival of the previous round's
        result (a one-thread flag wait on this stream): the last stage can no longer overwrite what is injected."""
        from distributed_llm_inference_b200 import _ffi

        r = self.rings[ring]
        dst = r.tok_inbox[m] if ring == "tokens" else r.inbox[m]
        if x.data_ptr() == dst.data_ptr():
            return
        if r.steps > 0:
            _ffi.check(_ffi.lib.dli_flag_wait(r.flags[m].data_ptr(), r.steps, r.state[m].data_ptr(), 10000,
                                              _ffi.current_stream(self.device)))
        dst.copy_(x.view(dst.shape) if x.numel() == dst.numel() else x, non_blocking=True)

    def _run_round_flags(self, r, generation_ids, inputs):
        """Every micro-batch's stage step, back to back on one stream; the cross-GPU order is carried by the flags the
        steps' own entry / exit kernels wait on and release."""
        first = self.rank == 0
        gs = [r.graphs[gid] for gid in generation_ids]
        if first and inputs is not None:
            for m, x in enumerate(inputs):
                self.inject(r.kind, m, x)
        for g in gs:
            g.before_replay()
        if r.round is not None and self.trace is None and tuple(generation_ids) == r.round[0]:
            r.round[1].replay()
        else:
            for g in gs:
                if self.trace is not None:
                    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
                    ev[0].record()
                g.graph.replay()
                if self.trace is not None:
                    ev[1].record()
                    self.trace.append(ev)
        for g in gs:
            g.after_replay()
        r.steps += 1
        self.launches_per_round = sum(g.launches for g in gs)
        if first:
            return [r.tok_inbox[m] if r.kind == "tokens" else r.inbox[m] for m in range(len(gs))]
        return None

    def step_one(self, ring: str, m: int) -> None:
        """Launch micro-batch ``m``'s stage step alone (host-fed rounds interleave host work between the steps); the
        caller finishes the round with ``end_round``."""
        g = self.rings[ring].graphs[self.rings[ring].gids[m]]
        g.before_replay()
        g.graph.replay()
        g.after_replay()

    def end_round(self, ring: str) -> None:
        self.rings[ring].steps += 1

    def wait_result(self, ring: str, m: int, stream) -> None:
        """Stage 0: make ``stream`` wait until the last stage's result of the LATEST launched round for micro-batch ``m``
        has arrived in this stage's inbox (one-thread kernel; time-out recorded in the step state)."""
        from distributed_llm_inference_b200 import _ffi

        r = self.rings[ring]
        _ffi.check(_ffi.lib.dli_flag_wait(r.flags[m].data_ptr(), r.steps, r.state[m].data_ptr(), 10000, stream.cuda_stream))

    def check_errors(self) -> None:
        """After a synchronise: raise if any in-kernel flag wait gave up (a producer never signalled)."""
        for kind, r in self.rings.items():
            if hasattr(r, "state") and bool(r.state[:, 2].any().item()):
                raise RuntimeError(f"stage {self.rank}: a hand-off wait of the {kind} ring timed out "
                                   f"(state {r.state.tolist()}, flags {r.flags[:, 0].tolist()})")

    def drain(self, n_micro_batches: int = 0) -> None:
        """Wait until everything launched so far -- including the ring closure into stage 0 -- has completed."""
        if str(self.device) != "cpu":
            for kind, r in self.rings.items():
                if self.rank == 0 and r.steps > 0 and hasattr(r, "flags"):
                    from distributed_llm_inference_b200 import _ffi

                    for m in range(len(r.gids)):
                        _ffi.check(_ffi.lib.dli_flag_wait(r.flags[m].data_ptr(), r.steps, r.state[m].data_ptr(), 10000,
                                                          _ffi.current_stream(self.device)))
            torch.cuda.synchronize(self.device)
