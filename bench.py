#!/usr/bin/env python
"""bench.py — decode throughput of a Llama span on B200 (BASELINE.json metric), with roofline and CPU baseline.

    python bench.py --gpus 1 --steps 64 --warmup 8             # this repo's arm
    python bench.py --impl reference --gpus 1 --steps 2 --warmup 1   # the reference's CPU path (oracle port)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...   # N-stage pipeline

Workload (config.workload): BASELINE.json configs[1] — Llama-2-7B-shaped bf16, one span of 32 layers on one B200,
batch 16, 512-token prompt, then KV-cached decode; one STEP = one decode step of the whole span for the batch
(16 tokens).  Weights are random-init of that architecture, prompts synthetic (no network).  For N > 1 the same
32 layers are cut into N pipeline stages (one process per GPU) with N micro-batches of 16 sequences in flight; each
stage's last kernel stores the hand-off tensor straight into the next stage's inbox over NVLink (NCCL send/recv is
the fallback).  Per-GPU work per step is the same at every N (N micro-batches x 32/N layers), the number of
sequences served grows with N: "weak" scaling.

Timing: W >= 3 warm-up steps; K timed steps bracketed by barrier + synchronize, CUDA events on the launching
stream, max over ranks.  Every step streams the span's 12.9 GB of weights (>> the 126 MB L2), so no explicit L2
flush is needed ("inputs larger than L2").
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

from distributed_llm_inference_b200.utils import synthetic as syn  # noqa: E402

METRIC = "decode_tokens_per_sec"
UNIT = "tokens/s"
MODEL = "7b"
BATCH = 16
PROMPT = 512
GEMM_CLASSES = ("gemm_qkv", "gemm_o", "gemm_gate_up", "gemm_down")


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "measured"
    return 6650.0, 1400.0, "fallback"


def measured_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the ncu --set full capture
    summarised under profiles/ (mean over the four GEMM launches of a layer); None if no capture is recorded."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("gemm_tcgen05_decode_bytes_per_launch")
    return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=1.0)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def decode_step_bytes(shape: syn.LlamaShape, n_layers: int, batch: int, ctx: int) -> dict:
    """ALGORITHMIC bytes of one decode stage-step: weights read once, K and V read once,
    K/V appended, span input + output."""
    h, i = shape.hidden_size, shape.intermediate_size
    H, KV, hd = shape.num_attention_heads, shape.num_key_value_heads, shape.head_dim
    per_gemm = {"gemm_qkv": 2 * h * (H + 2 * KV) * hd, "gemm_o": 2 * H * hd * h, "gemm_gate_up": 2 * 2 * h * i,
                "gemm_down": 2 * i * h}
    weights = n_layers * (sum(per_gemm.values()) + 2 * 2 * h)
    kv_read = n_layers * batch * ctx * 2 * KV * hd * 2
    kv_write = n_layers * batch * 2 * KV * hd * 2
    io = 2 * batch * h * 2
    return {"total": weights + kv_read + kv_write + io, "weights": weights, "kv_read": kv_read,
            "per_gemm_per_layer": per_gemm}


def build_span(shape: syn.LlamaShape, layer_ids, device, rope="float", max_positions: int = 8192):
    from distributed_llm_inference_b200.models.llama import LlamaBlock

    with torch.device(device):
        blk = LlamaBlock(shape.to_hf_config("eager", max_positions), list(layer_ids), rope=rope)
    for j, g in enumerate(layer_ids):   # random-init of the architecture, generated on the device
        w = syn.make_layer_weights(shape, g, device=device)
        blk.load_state_dict({f"layers.{j}.{n}.weight": t for n, t in w.items()}, strict=False)
        del w
    return blk


def gemm_chain_time(blk, shape: syn.LlamaShape, batch: int, dev, reps: int = 5):
    """Average launch duration of the dominant kernel as it runs in the step: the span's own 4 GEMMs per layer, over
    every layer's (distinct, 12.9 GB >> L2) weights, back to back on one stream through the C ABI (dli_gemm), CUDA
    events around the chain.  Unlike the per-class events inside dli_span_forward this keeps programmatic dependent
    launch active (an event record between two kernels serialises them)."""
    from distributed_llm_inference_b200 import _ffi

    h, i = shape.hidden_size, shape.intermediate_size
    qd = shape.num_attention_heads * shape.head_dim
    nqkv = qd + 2 * shape.num_key_value_heads * shape.head_dim
    xh = torch.randn(batch, h, device=dev).to(torch.bfloat16)
    xq = torch.randn(batch, qd, device=dev).to(torch.bfloat16)
    xi = torch.randn(batch, i, device=dev).to(torch.bfloat16)
    o_qkv = torch.zeros(batch, nqkv, device=dev)
    o_h = torch.zeros(batch, h, device=dev)
    o_act = torch.zeros(batch, i, device=dev, dtype=torch.bfloat16)
    st = _ffi.current_stream()

    def chain():
        for layer in blk.layers:
            a, m = layer.self_attn, layer.mlp
            for x, w, out, n, k, epi in ((xh, layer._wqkv, o_qkv, nqkv, h, _ffi.EPI_ADD_F32),
                                         (xq, a.o_proj.weight, o_h, h, qd, _ffi.EPI_ADD_F32),
                                         (xh, layer._wgu, o_act, i, h, _ffi.EPI_SILU_MUL_BF16),
                                         (xi, m.down_proj.weight, o_h, h, i, _ffi.EPI_ADD_F32)):
                _ffi.check(_ffi.lib.dli_gemm(x.data_ptr(), k, w.data_ptr(), out.data_ptr(), n, batch, n, k, epi, st))

    chain()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        chain()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * 4 * len(blk.layers))      # ms per launch


# ---------------------------------------------------------------------------------------------------------
# CPU baseline: the oracle port of the reference (torch fp32, all host threads) on a bounded sample
# ---------------------------------------------------------------------------------------------------------
def cpu_reference_decode(shape: syn.LlamaShape, n_layers_total: int, batch: int, ctx: int, budget_s: float = 20.0):
    """Times ONE layer of the exact shape (fp32 weights of all layers do not fit host RAM and a full pass takes
    minutes per token): prefill-free — the cache is filled with random K/V of length ctx — then decode steps until
    ~budget_s of CPU work; tokens/s is extrapolated as batch / (per-layer step time * n_layers_total)."""
    from oracle import span_oracle as so

    torch.set_num_threads(os.cpu_count() or 1)
    w = syn.make_layer_weights(shape, 0, dtype=torch.float32)
    oshape = so.SpanShape(shape.hidden_size, shape.intermediate_size, shape.num_attention_heads,
                          shape.num_key_value_heads, shape.head_dim, shape.rope_theta)
    span = so.OracleSpan(oshape, [so.LayerWeights(*(w[k] for k in syn.LINEARS + syn.NORMS))], "eager", "float")
    cache = so.SinkCacheOracle(1 << 20, 4)
    g = torch.Generator().manual_seed(5)
    kv = lambda: torch.randn(batch, shape.num_key_value_heads, ctx, shape.head_dim, generator=g)
    cache.update(kv(), kv(), 0, {"generation_id": "g", "cos": None, "sin": None})
    x = syn.make_hidden(shape, batch, 1, dtype=torch.float32)
    times, pos, t_start = [], ctx, time.perf_counter()
    with torch.no_grad():
        while True:
            t0 = time.perf_counter()
            span.forward("g", x, past_key_value=cache, cache_position=torch.tensor([pos]))
            times.append(time.perf_counter() - t0)
            pos += 1
            if len(times) >= 3 and (time.perf_counter() - t_start > budget_s or len(times) >= 64):
                break
    warm = sorted(times[1:])
    per_layer = warm[len(warm) // 2]
    return {"value": batch / (per_layer * n_layers_total), "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
            "sample": f"1 layer of {shape.name} shape, fp32 torch CPU oracle port of the reference, batch {batch}, "
                      f"ctx {ctx}, {len(times)} decode steps (median of warm), x{n_layers_total} layers extrapolated",
            "per_layer_ms": per_layer * 1e3}


# ---------------------------------------------------------------------------------------------------------
# this repo's arm
# ---------------------------------------------------------------------------------------------------------
def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def run_ours(args):
    import torch.distributed as dist

    from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
    from distributed_llm_inference_b200.parallel.pipeline import PipelineStage

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback for the span forward)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # the hand-off kernels must get SM slots while the stage's own kernels keep every SM full: high-priority stream
        opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
        dist.init_process_group("nccl", device_id=dev, pg_options=opts)
    shape = syn.SHAPES[MODEL]
    n_total = args.layers or shape.num_hidden_layers
    per = n_total // world
    layer_ids = list(range(rank * per, (rank + 1) * per if rank < world - 1 else n_total))
    K, W = args.steps, max(args.warmup, 3)
    n_mb = args.micro_batches or world   # micro-batches in flight (default: the pipeline depth)
    total_ctx = PROMPT + W + K + 8 + 32 + 8

    blk = build_span(shape, layer_ids, dev, max_positions=max(8192, total_ctx + 64))
    cache = PartialLlamaSinkCache(1 << 20, 4, page_tokens=args.page_tokens,
                                  capacity_tokens=n_mb * BATCH * (total_ctx + 16) + BATCH * 64 +
                                  (n_mb if world > 1 else 1) * BATCH * (PROMPT + 16) + (PROMPT + 128))
    stage = PipelineStage(blk, cache, rank, world, dev)
    gids = [f"mb{m}" for m in range(n_mb)]

    # ---- prefill (also timed: prefill tokens/s of the span/pipeline) ----
    # --prefill-chunk C: every prompt travels as ceil(PROMPT / C) chunks, so that stage s works on chunk c while stage
    # s + 1 works on chunk c - 1 (config C5: one 8k-token sequence over 4 or 8 stages)
    chunk = min(args.prefill_chunk or PROMPT, PROMPT)
    prompts = [syn.make_hidden(shape, BATCH, PROMPT, seed=99 + m, device=dev) for m in range(n_mb)]
    chunks = [(c0, [p[:, c0:c0 + chunk].contiguous() for p in prompts] if rank == 0 else None)
              for c0 in range(0, PROMPT, chunk)]

    def prefill_round(ids):
        for c0, ins in chunks:
            stage.run_round(ids, ins, BATCH, min(chunk, PROMPT - c0), c0, feed_back=False)

    if world > 1:   # untimed pass on throw-away generations: one-time descriptor / workspace / communicator set-up
        warm_ids = [f"pfw{m}" for m in range(n_mb)]
        prefill_round(warm_ids)
        torch.cuda.synchronize()
        for g in warm_ids:
            cache.close(g)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    prefill_round(gids)
    e1.record()
    torch.cuda.synchronize()
    prefill_ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    for g in gids:
        cache.reserve(g, BATCH, total_ctx)            # block tables stay fixed from here on

    # ---- warm prefill of the same chunk shape (tensor-core roofline), one GPU only ----
    prefill = None
    if world == 1:
        x_pf = prompts[0]
        times = []
        for rep in range(3):
            torch.cuda.synchronize()
            e0.record()
            blk("warm_prefill", x_pf, past_key_value=cache)
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1))
            if rep == 2:
                blk.profile(True)
                cache.close("warm_prefill")
                blk("warm_prefill", x_pf, past_key_value=cache)
                pf_prof = blk.profile_read()
                blk.profile(False)
            cache.close("warm_prefill")
        ms = min(times)
        h, i = shape.hidden_size, shape.intermediate_size
        H, KV, hd = shape.num_attention_heads, shape.num_key_value_heads, shape.head_dim
        T = PROMPT
        flops = len(layer_ids) * (2 * BATCH * T * (h * H * hd + 2 * h * KV * hd + H * hd * h + 3 * h * i)
                                  + BATCH * 2 * T * T * H * hd)
        _, tc_peak, pk = measured_peaks()
        prefill = {"tokens_per_sec": round(BATCH * T / (ms * 1e-3), 1), "ms": round(ms, 3),
                   "tflops": round(flops / (ms * 1e-3) / 1e12, 1), "peak_tflops": tc_peak,
                   "frac": round(flops / (ms * 1e-3) / 1e12 / tc_peak, 4),
                   "peak_source": f"MEASURED_PEAKS.json ({pk}, bf16_tflops_sustained)",
                   "flops": flops, "chunk": f"{BATCH} x {T} tokens, causal, warm (best of 3)",
                   "kernel_ms": {c: round(v[0], 3) for c, v in pf_prof.items()}}

    # ---- decode: warm-up, then K timed steps ----
    # one GPU: the whole stage step is one CUDA graph (LlamaBlock.graphed_decode); pipeline: one graph per stage and
    # micro-batch with peer-store hand-off, then (after warm-up) one graph per stage and round
    xs = [syn.make_hidden(shape, BATCH, 1, seed=7 + m, device=dev) for m in range(n_mb)] if rank == 0 else None
    pos = PROMPT
    graph = None
    if not args.no_graph:
        stage.enable_graphs(gids, total_ctx, batch=BATCH, peer_store=not args.nccl_handoff)
        graph = stage.graphs[gids[0]]

    def decode_round(xs, pos):
        return stage.run_round(gids, xs, BATCH, 1, pos, feed_back=True)

    for _ in range(W):
        xs = decode_round(xs, pos)
        pos += 1
    round_graph = False
    if world > 1 and graph is not None and stage.handoff == "peer_store" and not args.no_round_graph:
        round_graph = stage.capture_round(gids)     # steady-state rounds: one graph launch per round
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    lat = []
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    evs[0].record()
    for k in range(K):
        xs = decode_round(xs, pos)
        pos += 1
        evs[k + 1].record()
    torch.cuda.synchronize()
    if os.environ.get("BENCH_TRACE"):
        stage.trace = []
        for _ in range(8):
            xs = decode_round(xs, pos)
            pos += 1
        torch.cuda.synchronize()
        tr = stage.trace[n_mb:]
        print(f"[trace rank {rank}] recv-wait {sum(e[0].elapsed_time(e[1]) for e in tr) / len(tr):.3f} ms, step "
              f"{sum(e[1].elapsed_time(e[2]) for e in tr) / len(tr):.3f} ms per micro-batch; between steps "
              f"{sum(a[2].elapsed_time(b[0]) for a, b in zip(tr[:-1], tr[1:])) / (len(tr) - 1):.3f} ms", flush=True)
        stage.trace = None
    clocks = sampler.finish()
    if world > 1:
        dist.barrier()
    total_ms = torch.tensor([evs[0].elapsed_time(evs[K])], device=dev)
    lat = sorted(evs[k].elapsed_time(evs[k + 1]) for k in range(K))
    launches = stage.launches_per_round * K
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(prefill_ms, op=dist.ReduceOp.MAX)
    ms_per_step = total_ms.item() / K
    tokens_per_step = BATCH * n_mb
    value = tokens_per_step / (ms_per_step * 1e-3)
    ctx_mid = PROMPT + W + K // 2

    # ---- instrumented pass: per-kernel-class time with CUDA events inside dli_span_forward ----
    stage.drain(n_mb)
    saved_graphs, stage.graphs = stage.graphs, {}          # events cannot be recorded inside a replayed graph
    blk.profile(True)
    n_prof = min(K, 8)
    for _ in range(n_prof):
        xs = stage.run_round(gids, xs, BATCH, 1, pos, feed_back=True)
        pos += 1
    prof = blk.profile_read()
    blk.profile(False)
    stage.graphs = saved_graphs
    hbm_peak, tc_peak, peak_kind = measured_peaks()
    sb = decode_step_bytes(shape, len(layer_ids), BATCH, ctx_mid)
    gemm_ms = sum(prof[c][0] for c in GEMM_CLASSES)
    gemm_launches = sum(prof[c][1] for c in GEMM_CLASSES)
    step_prof_ms = sum(v[0] for v in prof.values()) / (n_prof * n_mb)
    gemm_bytes_per_launch = sum(sb["per_gemm_per_layer"].values()) / 4.0
    gemm_ms_serialised = gemm_ms / max(1, gemm_launches)
    gemm_ms_per_launch = gemm_chain_time(blk, shape, BATCH, dev)
    achieved = gemm_bytes_per_launch / (gemm_ms_per_launch * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "gemm_tcgen05_kernel (swap-AB split-K weight streaming; 4 launches/layer)",
                "achieved": round(achieved, 1), "peak": hbm_peak, "unit": "GB/s", "frac": round(achieved / hbm_peak, 4),
                "traffic": measured_traffic(), "peak_source": f"MEASURED_PEAKS.json ({peak_kind}, hbm_gbs)",
                "bytes_per_launch": gemm_bytes_per_launch, "ms_per_launch": round(gemm_ms_per_launch, 5),
                "how": "CUDA events around the span's own 4 GEMMs/layer x all layers launched back to back through "
                       "dli_gemm (programmatic dependent launch active, as in the step)",
                "ms_per_launch_serialised": round(gemm_ms_serialised, 5),
                "share_of_step": round(gemm_ms / max(1e-9, sum(v[0] for v in prof.values())), 4)}
    stage_gbs = sb["total"] / (ms_per_step / n_mb * 1e-3) / 1e9
    breakdown = {c: round(v[0] / (n_prof * n_mb), 4) for c, v in prof.items()}

    # ---- e2e: host buffers through the public API, H2D + D2H inside the timed region ----
    e2e = None
    if world == 1:
        hin = [torch.empty(BATCH, 1, shape.hidden_size, dtype=torch.bfloat16).pin_memory() for _ in range(2)]
        hout = torch.empty(BATCH, 1, shape.hidden_size, dtype=torch.bfloat16).pin_memory()
        for t in hin:
            t.copy_(syn.make_hidden(shape, BATCH, 1, seed=3))
        n_e2e = min(K, 32)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for k in range(n_e2e):
            x = hin[k & 1].to(dev, non_blocking=True)
            if graph is not None:
                y = graph.step(x)
            else:
                y = blk(gids[0], x, past_key_value=cache, cache_position=torch.tensor([pos], device=dev))[0]
            hout.copy_(y, non_blocking=True)
            torch.cuda.synchronize()               # the caller needs the result before the next token
            pos += 1
        dt = time.perf_counter() - t0
        nbytes = BATCH * shape.hidden_size * 2
        e2e = {"value": round(BATCH * n_e2e / dt, 1), "unit": UNIT, "h2d_bytes_per_step": nbytes,
               "d2h_bytes_per_step": nbytes, "steps": n_e2e,
               "how": ("LlamaBlock.graphed_decode(...).step" if graph is not None else "LlamaBlock.forward") +
                      " per step with pinned-host input copied H2D and output copied D2H, synchronised every step "
                      "(wall clock)"}

    elif graph is not None and stage.handoff == "peer_store":
        # pipeline: every step, stage 0 copies each micro-batch's input from pinned host memory into the head of the pipe,
        # the round runs through all stages, the last stage's peer stores land back on stage 0 and are copied to the
        # host; the next step starts only then (a client needs token t before it can feed token t + 1)
        stage.drain(n_mb)
        n_e2e = min(K, 16)
        g0 = [stage.graphs[g] for g in gids]
        if rank == 0:
            hin = [torch.empty(BATCH, 1, shape.hidden_size, dtype=torch.bfloat16).pin_memory() for _ in range(n_mb)]
            hout = [torch.empty(BATCH, 1, shape.hidden_size, dtype=torch.bfloat16).pin_memory() for _ in range(n_mb)]
            for m, t in enumerate(hin):
                t.copy_(syn.make_hidden(shape, BATCH, 1, seed=3 + m))
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        for k in range(n_e2e):
            if rank == 0:
                for m in range(n_mb):
                    g0[m].x.copy_(hin[m], non_blocking=True)
                stage.run_round(gids, [g.x for g in g0], BATCH, 1, pos, feed_back=True)
                stage.drain(n_mb)
                for m in range(n_mb):
                    hout[m].copy_(g0[m].x, non_blocking=True)
                torch.cuda.synchronize()
            else:
                stage.run_round(gids, None, BATCH, 1, pos, feed_back=True)
            pos += 1
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], device=dev)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        nbytes = n_mb * BATCH * shape.hidden_size * 2
        e2e = {"value": round(n_mb * BATCH * n_e2e / dt.item(), 1), "unit": UNIT, "h2d_bytes_per_step": nbytes,
               "d2h_bytes_per_step": nbytes, "steps": n_e2e,
               "how": f"PipelineStage.run_round over {world} stages, {n_mb} micro-batches per step: pinned-host inputs copied "
                      "H2D on stage 0, outputs of the last stage copied D2H on stage 0, synchronised every step (wall "
                      "clock, max over ranks); steps do not overlap, so the pipe fills and drains every step"}

    # ---- configs[1] also names batch 1: per-token latency of the same span with ONE sequence (one GPU only) ----
    batch1 = None
    if world == 1 and graph is not None and BATCH != 1:
        x1 = syn.make_hidden(shape, 1, PROMPT, seed=123, device=dev)
        blk("b1", x1, past_key_value=cache)
        n1 = min(K, 32)
        g1 = blk.graphed_decode("b1", cache, PROMPT + n1 + 8)
        xt1 = syn.make_hidden(shape, 1, 1, seed=124, device=dev)
        for _ in range(3):
            g1.step(xt1)
        ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(n1 + 1)]
        torch.cuda.synchronize()
        ev1[0].record()
        for k in range(n1):
            g1.step()
            ev1[k + 1].record()
        torch.cuda.synchronize()
        lat1 = sorted(ev1[k].elapsed_time(ev1[k + 1]) for k in range(n1))
        ms1 = ev1[0].elapsed_time(ev1[n1]) / n1
        sb1 = decode_step_bytes(shape, len(layer_ids), 1, PROMPT + 3 + n1 // 2)
        batch1 = {"tokens_per_sec": round(1e3 / ms1, 1), "ms_per_step": round(ms1, 4), "p50_ms_per_token": round(lat1[n1 // 2], 4),
                  "steps": n1, "stage_roofline_frac": round(sb1["total"] / (ms1 * 1e-3) / 1e9 / hbm_peak, 4),
                  "bytes_per_step": sb1["total"]}
        cache.close("b1")

    if rank == 0:
        cpu = (cpu_reference_decode(shape, n_total, BATCH, PROMPT, budget_s=args.cpu_budget)
               if (world == 1 and args.cpu_budget > 0) else None)
        pf_pipe = None
        if world > 1:   # whole-pipeline prefill: all stages' FLOPs over the max-over-ranks time, against N x the peak
            h_, i_ = shape.hidden_size, shape.intermediate_size
            H_, KV_, hd_ = shape.num_attention_heads, shape.num_key_value_heads, shape.head_dim
            fl = n_mb * n_total * (2 * BATCH * PROMPT * (h_ * H_ * hd_ + 2 * h_ * KV_ * hd_ + H_ * hd_ * h_ + 3 * h_ * i_)
                                   + BATCH * 2 * PROMPT * PROMPT * H_ * hd_)
            _, tc_peak_, _ = measured_peaks()
            pf_ms = prefill_ms.item()
            pf_pipe = {"ms": round(pf_ms, 3), "chunk_tokens": chunk, "chunks": len(chunks), "sequences": n_mb * BATCH,
                       "tflops": round(fl / (pf_ms * 1e-3) / 1e12, 1),
                       "frac_of_n_gpu_peak": round(fl / (pf_ms * 1e-3) / 1e12 / (tc_peak_ * world), 4),
                       "fill_drain_bound": round(len(chunks) * n_mb / (len(chunks) * n_mb + world - 1), 4)}
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"{shape.name}-shaped span ({n_total} layers) KV-cached decode, batch {BATCH} x "
                                   f"{n_mb} micro-batch(es) in flight, {PROMPT}-token prompt, context "
                                   f"{PROMPT + W}..{PROMPT + W + K}; random-init weights, synthetic hidden states",
                       "global_batch": tokens_per_step, "seq_len": PROMPT, "parallelism": f"pp{world}",
                       "l2": "inputs larger than L2 (12.9 GB of weights streamed per step)", "page_tokens": args.page_tokens,
                       "launch": ("cuda_graph_per_round" if round_graph else "cuda_graph") if graph is not None else "eager",
                       "round_graph_error": stage.round_error,
                       "handoff": stage.handoff, "handoff_fallback_reason": stage.handoff_error},
            "p50_ms_per_token": round(lat[len(lat) // 2] / n_mb, 4),
            "prefill_tokens_per_sec": (prefill["tokens_per_sec"] if prefill is not None else
                                       round(n_mb * BATCH * PROMPT / (prefill_ms.item() * 1e-3), 1)),
            "prefill_first_call_tokens_per_sec": round(n_mb * BATCH * PROMPT / (prefill_ms.item() * 1e-3), 1),
            "prefill_note": ("pipelined over the stages in chunks of %d tokens, warm" % chunk) if world > 1 else
                            "warm chunk (details in `prefill`); the first call also pays one-time descriptor/workspace set-up",
            "prefill_pipeline": pf_pipe, "prefill": prefill,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roofline,
            "stage_roofline": {"bound": "hbm", "achieved": round(stage_gbs, 1), "peak": hbm_peak, "unit": "GB/s",
                               "frac": round(stage_gbs / hbm_peak, 4), "bytes_per_stage_step": sb["total"],
                               "note": "whole decode stage-step (all kernels + gaps), SURVEY 8d algorithmic bytes"},
            "kernel_ms_per_step": breakdown, "kernel_ms_per_step_sum": round(step_prof_ms, 4),
            "e2e": e2e, "cpu_baseline": cpu, "decode_batch1": batch1,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    """The reference arm: the reference's own CPU implementation of the path.  The reference is Python on top of
    transformers 4.43.3 + hivemind, neither installable here, and cannot travel to the GPU box; what is timed is
    the oracle port (oracle/span_oracle.py, pinned bit-exactly to the reference on this container) on the host
    cores, on this arm's config, metric and unit.  Each step is a bounded sample: one layer of the shape, one
    decode step, extrapolated x n_layers."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    shape = syn.SHAPES[MODEL]
    K, W = args.steps, args.warmup
    budget = min(60.0, max(5.0, 4.0 * (K + W)))
    cpu = cpu_reference_decode(shape, shape.num_hidden_layers, BATCH, PROMPT, budget_s=budget)
    line = {"impl": "reference", "metric": METRIC, "value": round(cpu["value"], 3), "unit": UNIT, "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": round(BATCH / cpu["value"] * 1e3, 2), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"Llama-2-7B-shaped span (32 layers) KV-cached decode, batch {BATCH}, "
                                   f"{PROMPT}-token context, CPU", "global_batch": BATCH, "seq_len": PROMPT,
                       "parallelism": "cpu"},
            "cpu_baseline": cpu,
            "e2e": {"value": round(cpu["value"], 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--batch", type=int, default=None, help="override the batch (experiments; the default is the config's 16)")
    ap.add_argument("--model", default=None, choices=["7b", "13b", "70b", "tiny"], help="shape (default: configs[1], 7b)")
    ap.add_argument("--page-tokens", type=int, default=16, help="tokens per KV page")
    ap.add_argument("--prompt", type=int, default=None, help="prompt length (default 512, the config's)")
    ap.add_argument("--layers", type=int, default=None, help="total layers over all stages (default: the model's)")
    ap.add_argument("--micro-batches", type=int, default=None, help="micro-batches in flight (default: number of stages)")
    ap.add_argument("--prefill-chunk", type=int, default=None, help="prefill the prompt in chunks of this many tokens")
    ap.add_argument("--nccl-handoff", action="store_true", help="hand off by NCCL isend/irecv instead of peer stores")
    ap.add_argument("--no-graph", action="store_true", help="decode with eager launches instead of the CUDA graph")
    ap.add_argument("--no-round-graph", action="store_true", help="pipeline: one graph per micro-batch instead of per round")
    args = ap.parse_args()
    global BATCH, MODEL, PROMPT
    if args.prompt:
        PROMPT = args.prompt
    if args.batch:
        BATCH = args.batch
    if args.model:
        MODEL = args.model
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
