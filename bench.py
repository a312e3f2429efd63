#!/usr/bin/env python
"""bench.py — decode throughput of a Llama span on B200 (BASELINE.json metric), with roofline and CPU baseline.

    python bench.py --gpus 1 --steps 64 --warmup 8             # this repo's arm
    python bench.py --impl reference --gpus 1 --steps 2 --warmup 1   # the reference's own CPU path (oracle/_ref)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...   # N-stage pipeline

Workload (config.workload): BASELINE.json configs[1] — Llama-2-7B-shaped bf16, one span of 32 layers on one B200,
batch 16, 512-token prompt, then KV-cached decode; one STEP = one decode step of the whole span for the batch
(16 tokens).  Weights are random-init of that architecture, prompts synthetic (no network).  For N > 1 the same
32 layers are cut into N pipeline stages (one process per GPU) with N micro-batches of 16 sequences in flight; a
stage step's own entry / exit kernels do the hand-off over NVLink peer memory (flag wait, peer store, flag release).
Per-GPU work per step is the same at every N (N micro-batches x 32/N layers), the number of sequences served grows
with N: "weak" scaling.

What an N > 1 line carries besides the 7B weak-scaling value:
  pipeline_parity  before any timing, a prompt and 4 decode rounds go through the N stages (peer-store hand-off, round
                   graph included) and are compared on rank 0 with the same 32 layers run as ONE span, whose first two
                   layers are in turn compared with the CPU oracle; the token ring (embedding -> stages -> head ->
                   argmax) is compared with the sequential client.  A mismatch makes the run exit non-zero.
  configs          BASELINE.json configs 3-5 measured in the same process: N = 2 -> 13B 2 x 20 layers, batch 8;
                   N = 8 -> 70B 8 x 10 layers, decode with 32 sequences as (i) 1 micro-batch x 32, (ii) 8 x 4,
                   (iii) 8 x 32 in flight, and the 1 x 8192-token prefill in chunks; N = 4 -> that prefill on 4 stages.
  e2e              the same decode through the sequential client closed over the stages (token ring): every step the
                   host reads each micro-batch's token ids (D2H) and feeds them back (H2D) while the other micro-batches
                   are in flight.

Timing: W >= 3 warm-up steps; K timed steps bracketed by barrier + synchronize, CUDA events on the launching
stream, max over ranks.  Every step streams the span's 12.9 GB of weights (>> the 126 MB L2), so no explicit L2
flush is needed ("inputs larger than L2").
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

from distributed_llm_inference_b200.parallel.pipeline import PREFILL_CHUNK_TOKENS  # noqa: E402
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


def measured_traffic(shape_name: str, batch: int):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel from the ncu --set full capture
    summarised under profiles/traffic.json (mean over the four GEMM launches of a layer), for the SHAPE the capture was
    taken on; None for any other shape (no capture, no claim)."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            t = json.load(f)
        if t.get("shape", "llama2-7b") == shape_name and int(t.get("batch", 16)) == batch:
            return t.get("gemm_tcgen05_decode_bytes_per_launch")
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


def decode_step_bytes(shape: syn.LlamaShape, n_layers: int, batch: int, ctx: int, head: bool = False, w8: bool = False) -> dict:
    """ALGORITHMIC bytes of one decode stage-step: weights read once, K and V read once,
    K/V appended, span input + output; with ``head`` also the [vocab, h] head read once and the logits written and read."""
    h, i = shape.hidden_size, shape.intermediate_size
    H, KV, hd = shape.num_attention_heads, shape.num_key_value_heads, shape.head_dim
    per_gemm = {"gemm_qkv": 2 * h * (H + 2 * KV) * hd, "gemm_o": 2 * H * hd * h, "gemm_gate_up": 2 * 2 * h * i,
                "gemm_down": 2 * i * h}
    if w8:   # int8 weights + one fp32 scale per output row
        rows = (H + 2 * KV) * hd + h + 2 * i + h
        per_gemm = {k: v // 2 for k, v in per_gemm.items()}
        weights = n_layers * (sum(per_gemm.values()) + 4 * rows + 2 * 2 * h)
    else:
        weights = n_layers * (sum(per_gemm.values()) + 2 * 2 * h)
    kv_read = n_layers * batch * ctx * 2 * KV * hd * 2
    kv_write = n_layers * batch * 2 * KV * hd * 2
    io = 2 * batch * h * 2
    head_bytes = (2 * shape.vocab_size * h + 2 * batch * shape.vocab_size * 4) if head else 0
    return {"total": weights + kv_read + kv_write + io + head_bytes, "weights": weights, "kv_read": kv_read,
            "head": head_bytes, "per_gemm_per_layer": per_gemm}


def prefill_flops(shape: syn.LlamaShape, n_layers: int, batch: int, T: int) -> int:
    h, i = shape.hidden_size, shape.intermediate_size
    H, KV, hd = shape.num_attention_heads, shape.num_key_value_heads, shape.head_dim
    return n_layers * (2 * batch * T * (h * H * hd + 2 * h * KV * hd + H * hd * h + 3 * h * i) + batch * 2 * T * T * H * hd)


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
# CPU baseline: the reference's own files (run-time copy under oracle/_ref) or, without it, the oracle port
# ---------------------------------------------------------------------------------------------------------
REF_COPY = os.path.join(ROOT, "oracle", "_ref")


def _reference_available() -> bool:
    return os.path.isdir(os.path.join(REF_COPY, "distributed_llm_inference", "models", "llama"))


def _cpu_span(shape: syn.LlamaShape, layer_ids, kind: str):
    """-> (span callable, cache factory): the reference's LlamaBlock through oracle/ref_shim.py (kind "reference") or
    the oracle restatement (kind "port"), fp32 on the host."""
    W = [syn.make_layer_weights(shape, g, dtype=torch.float32) for g in layer_ids]
    if kind == "reference":
        os.environ["DLI_REFERENCE_ROOT"] = REF_COPY
        from oracle import ref_shim

        cfg = ref_shim.make_config(shape.hidden_size, shape.intermediate_size, shape.num_attention_heads,
                                   shape.num_key_value_heads, "eager", shape.rope_theta)
        blk = ref_shim.make_block(cfg, len(W), rope="float")
        ref_shim.load_layer_weights(blk, W)

        def span(gid, x, cache, cp):
            return ref_shim.block_forward(blk, gid, x, cache, cache_position=cp)[0]

        return span, lambda: ref_shim.make_cache(1 << 20, 4)
    from oracle import span_oracle as so

    oshape = so.SpanShape(shape.hidden_size, shape.intermediate_size, shape.num_attention_heads,
                          shape.num_key_value_heads, shape.head_dim, shape.rope_theta)
    osp = so.OracleSpan(oshape, [so.LayerWeights(*(w[k] for k in syn.LINEARS + syn.NORMS)) for w in W], "eager", "float")

    def span(gid, x, cache, cp):
        return osp.forward(gid, x, past_key_value=cache, cache_position=cp)[0]

    return span, lambda: so.SinkCacheOracle(1 << 20, 4)


def cpu_reference_decode(shape: syn.LlamaShape, n_layers_total: int, batch: int, ctx: int, budget_s: float = 20.0,
                         real_prefill: bool = False):
    """Times ONE layer of the exact shape on the host cores (fp32 weights of all layers do not fit host RAM and a full
    pass takes minutes per token): the layer's cache is filled by a real ``ctx``-token prefill (``real_prefill``; the
    reference arm) or with random K/V of that length (the quick cpu_baseline leg of the GPU arm), then decode steps run
    until ~budget_s of CPU work; tokens/s is EXTRAPOLATED as batch / (per-layer step time * n_layers_total) and the
    record says so."""
    torch.set_num_threads(os.cpu_count() or 1)
    kind = "reference" if _reference_available() else "port"
    span, make_cache = _cpu_span(shape, [0], kind)
    cache = make_cache()
    t_pf = None
    with torch.no_grad():
        if real_prefill:
            xp = syn.make_hidden(shape, batch, ctx, dtype=torch.float32, seed=9)
            t0 = time.perf_counter()
            span("g", xp, cache, None)
            t_pf = time.perf_counter() - t0
        else:
            g = torch.Generator().manual_seed(5)
            kv = lambda: torch.randn(batch, shape.num_key_value_heads, ctx, shape.head_dim, generator=g)
            cache.update(kv(), kv(), 0, {"generation_id": "g", "cos": None, "sin": None})
        x = syn.make_hidden(shape, batch, 1, dtype=torch.float32)
        times, pos, t_start = [], ctx, time.perf_counter()
        while True:
            t0 = time.perf_counter()
            span("g", x, cache, torch.tensor([pos]))
            times.append(time.perf_counter() - t0)
            pos += 1
            if len(times) >= 3 and (time.perf_counter() - t_start > budget_s or len(times) >= 64):
                break
    warm = sorted(times[1:])
    per_layer = warm[len(warm) // 2]
    what = ("the reference's own LlamaBlock / PartialLlamaSinkCache (files copied at build time to oracle/_ref, imported "
            "through oracle/ref_shim.py)" if kind == "reference" else
            "fp32 torch CPU oracle port of the reference (oracle/_ref absent)")
    out = {"value": batch / (per_layer * n_layers_total), "unit": UNIT, "cores": os.cpu_count(), "kind": kind,
           "extrapolated": True,
           "sample": f"1 layer of {shape.name} shape, {what}, fp32, batch {batch}, ctx {ctx} "
                     f"({'real prefill' if real_prefill else 'cache filled with random K/V'}), {len(times)} decode steps "
                     f"(median of warm), x{n_layers_total} layers EXTRAPOLATED",
           "per_layer_ms": per_layer * 1e3}
    if t_pf is not None:
        out["prefill_per_layer_s"] = round(t_pf, 3)
        out["prefill_tokens_per_sec_extrapolated"] = round(batch * ctx / (t_pf * n_layers_total), 2)
    return out


def cpu_reference_config1(budget_steps: int = 32):
    """BASELINE.json configs[0] measured IN FULL on the host cores: tiny Llama (12 layers, h = 768) as 2 spans x 6
    layers, batch 1, 128-token prompt, 32 greedy decode steps through the sequential client -- the reference's own
    files when oracle/_ref is present, else the oracle port."""
    from distributed_llm_inference_b200.client import SequentialClient

    torch.set_num_threads(os.cpu_count() or 1)
    kind = "reference" if _reference_available() else "port"
    shp = syn.SHAPES["tiny"]
    spans, caches = [], []
    for a, b in ((0, 6), (6, 12)):
        span, make_cache = _cpu_span(shp, list(range(a, b)), kind)
        spans.append(lambda gid, h, past_key_value=None, cache_position=None, _s=span: (_s(gid, h, past_key_value, cache_position),))
        caches.append(make_cache())
    cw = syn.make_client_weights(shp, dtype=torch.float32)
    client = SequentialClient(cw["embed_tokens"], cw["norm"], cw["lm_head"], spans, caches)
    prompt = syn.make_prompt(shp, 1, 128)
    with torch.no_grad():
        t0 = time.perf_counter()
        logits = client.step("c1", prompt, 0)
        t_pf = time.perf_counter() - t0
        lat, pos = [], 128
        for _ in range(budget_steps):
            nxt = logits.argmax(-1)
            t0 = time.perf_counter()
            logits = client.step("c1", nxt[:, None], pos)
            lat.append(time.perf_counter() - t0)
            pos += 1
    lat.sort()
    return {"kind": kind, "cores": os.cpu_count(), "prefill_tokens_per_sec": round(128 / t_pf, 1),
            "decode_tokens_per_sec": round(len(lat) / sum(lat), 2), "p50_ms_per_token": round(lat[len(lat) // 2] * 1e3, 3),
            "workload": "configs[0]: tiny Llama, 2 spans x 6 layers, batch 1, 128-token prompt (first call), "
                        f"{budget_steps} greedy steps, fp32, measured in full"}


# ---------------------------------------------------------------------------------------------------------
# this repo's arm
# ---------------------------------------------------------------------------------------------------------
def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


class Ctx:
    def __init__(self, args):
        import torch.distributed as dist

        self.args = args
        self.rank, self.world, self.local = dist_env()
        if not torch.cuda.is_available():
            raise RuntimeError("bench.py needs a CUDA device (no CPU fallback for the span forward)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = dist
        if self.world > 1:
            opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
            dist.init_process_group("nccl", device_id=self.dev, pg_options=opts)

    def barrier(self):
        torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def max_over_ranks(self, v: float) -> float:
        t = torch.tensor([v], device=self.dev, dtype=torch.float64)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.item()

    def all_ok(self, ok: bool) -> bool:
        t = torch.tensor([1 if ok else 0], device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return t.item() == 1


def stage_layers(n_total: int, rank: int, world: int):
    per = n_total // world
    return list(range(rank * per, (rank + 1) * per if rank < world - 1 else n_total))


def free_cuda():
    gc.collect()
    torch.cuda.empty_cache()


def run_workload(ctx: Ctx, model: str, batch: int, prompt: int, n_mb: int, layers_total, K: int, W: int,
                 prefill_chunk=None, full: bool = False, want_e2e: bool = False, page_tokens: int = 16, w8: bool = False, a8: bool = False):
    """One decode measurement of ``model`` cut into ``ctx.world`` stages: prefill (timed), warm-up, K timed rounds of
    ``n_mb`` micro-batches x ``batch`` sequences over the hidden-state ring.  ``full``: also the instrumented per-class
    pass, the dominant-kernel roofline, batch-1 latency, warm prefill (one GPU).  ``want_e2e``: the token ring
    through the sequential client with host-fed tokens.  Returns a dict (identical on every rank where it matters)."""
    from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
    from distributed_llm_inference_b200.parallel.pipeline import PipelineStage

    args, rank, world, dev, dist = ctx.args, ctx.rank, ctx.world, ctx.dev, ctx.dist
    shape = syn.SHAPES[model]
    n_total = layers_total or shape.num_hidden_layers
    layer_ids = stage_layers(n_total, rank, world)
    first, last = rank == 0, rank == world - 1
    n_e2e = min(K, 16) if want_e2e else 0
    n_prof = min(K, 8) if full else 0
    total_ctx = prompt + W + K + n_prof + n_e2e + 3 + 48
    hbm_peak, tc_peak, peak_kind = measured_peaks()

    blk = build_span(shape, layer_ids, dev, max_positions=max(8192, total_ctx + 64))
    if w8:
        blk.quantize_w8(activations="int8" if a8 else "bf16", threshold=args.a8_threshold)   # row-wise int8 weights (convert_to_optimized_block(quantize=True))
        free_cuda()
    cache = PartialLlamaSinkCache(1 << 20, 4, page_tokens=page_tokens,
                                  capacity_tokens=n_mb * batch * (total_ctx + 16) + batch * 192 +
                                  (n_mb if world > 1 else 1) * batch * (prompt + 16) + (prompt + 128))
    client_w = None
    if want_e2e:   # client tensors of the token ring (embedding on stage 0, final norm + head on the last stage)
        cw = syn.make_decidable_client_weights(shape) if (first or last) else {}
        client_w = {}
        if first:
            client_w["embed_tokens"] = cw["embed_tokens"].to(dev)
        if last:
            client_w.update(norm=cw["norm"].to(dev), lm_head=cw["lm_head"].to(dev), norm_eps=1e-6)
        del cw
    stage = PipelineStage(blk, cache, rank, world, dev, client=client_w)
    stage.chain_steps = not args.no_chain
    gids = [f"mb{m}" for m in range(n_mb)]

    # ---- prefill (also timed: prefill tokens/s of the span/pipeline) ----
    # prefill_chunk C: every prompt travels as ceil(prompt / C) chunks, so that stage s works on chunk c while stage
    # s + 1 works on chunk c - 1 (config C5: one 8k-token sequence over 4 or 8 stages)
    chunk = min(prefill_chunk or prompt, prompt)
    prompts = [syn.make_hidden(shape, batch, prompt, seed=99 + m, device=dev) for m in range(n_mb)] if first else None

    def prefill_round(ids):
        stage.prefill(ids, prompts, batch, prompt, chunk_tokens=chunk)

    if world > 1:   # untimed pass on throw-away generations: one-time descriptor / workspace / communicator set-up
        warm_ids = [f"pfw{m}" for m in range(n_mb)]
        prefill_round(warm_ids)
        torch.cuda.synchronize()
        for g in warm_ids:
            cache.close(g)
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    prefill_round(gids)
    e1.record()
    torch.cuda.synchronize()
    prefill_ms = ctx.max_over_ranks(e0.elapsed_time(e1))
    for g in gids:
        cache.reserve(g, batch, total_ctx)            # block tables stay fixed from here on

    # ---- warm prefill of the same chunk shape (tensor-core roofline), one GPU only ----
    prefill = None
    if world == 1 and full:
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
        flops = prefill_flops(shape, len(layer_ids), batch, prompt)
        prefill = {"tokens_per_sec": round(batch * prompt / (ms * 1e-3), 1), "ms": round(ms, 3),
                   "tflops": round(flops / (ms * 1e-3) / 1e12, 1), "peak_tflops": tc_peak,
                   "frac": round(flops / (ms * 1e-3) / 1e12 / tc_peak, 4),
                   "peak_source": f"MEASURED_PEAKS.json ({peak_kind}, bf16_tflops_sustained)",
                   "flops": flops, "chunk": f"{batch} x {prompt} tokens, causal, warm (best of 3)",
                   "kernel_ms": {c: round(v[0], 3) for c, v in pf_prof.items()}}
    del prompts

    # ---- decode: warm-up, then K timed rounds over the hidden-state ring ----
    # one GPU: the whole stage step is one CUDA graph (LlamaBlock.graphed_decode); pipeline: the stage step's own entry /
    # exit kernels carry the hand-off (flag wait, peer store, flag release), one graph per stage and round
    xs = [syn.make_hidden(shape, batch, 1, seed=7 + m, device=dev) for m in range(n_mb)] if first else None
    pos = prompt
    graph = None
    if world == 1:
        graph = blk.graphed_decode(gids[0], cache, total_ctx) if not args.no_graph else None
        graphs1 = [graph] + [blk.graphed_decode(g, cache, total_ctx) for g in gids[1:]] if graph is not None else None

        def decode_round(inputs):
            outs = []
            for m, gid in enumerate(gids):
                if graphs1 is not None:
                    outs.append(graphs1[m].step(inputs[m]))
                else:
                    outs.append(blk(gid, inputs[m], past_key_value=cache,
                                    cache_position=torch.tensor([cache.generation(gid).length], device=dev))[0])
            stage.launches_per_round = sum(g.launches for g in graphs1) if graphs1 is not None else blk.last_launch_count * n_mb
            return outs
    else:
        if not args.no_graph:
            stage.enable_graphs(gids, total_ctx, batch=batch, peer_store=not args.nccl_handoff, ring="hidden")

        def decode_round(inputs):
            return stage.run_round(gids, inputs, batch, 1, pos, feed_back=True)

    for w_ in range(W):
        xs = decode_round(xs)
        if world > 1 and stage.handoff == "peer_store":
            xs = None                                   # the ring feeds itself from its inbox
        pos += 1
    round_graph = False
    if world > 1 and stage.handoff == "peer_store" and not args.no_round_graph:
        round_graph = stage.capture_round(gids)     # steady-state rounds: one graph launch per round
    # every stage's step timed WITHOUT hand-off (nothing to wait for), all stages at the same time, each on its own GPU:
    # a pipeline with as many micro-batches as stages runs at the pace of its slowest stage, so this is the bound the
    # round time is read against.  The ~0.3 s of back-to-back steps also bring every GPU to its steady clocks (the GPUs
    # idle through the set-up; a few warm-up rounds of a few ms each do not do that).
    solo = None
    if not args.no_graph:
        ctx.barrier()
        xs_solo = syn.make_hidden(shape, batch, prompt, seed=77, device=dev)
        blk("solo", xs_solo, past_key_value=cache)
        del xs_solo
        gs = blk.graphed_decode("solo", cache, prompt + 120)
        for _ in range(100):
            gs.step()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(12):
            gs.step()
        e1.record()
        torch.cuda.synchronize()
        mine = torch.tensor([e0.elapsed_time(e1) / 12], device=dev)
        allms = [torch.zeros_like(mine) for _ in range(world)]
        if world > 1:
            dist.all_gather(allms, mine)
        else:
            allms = [mine]
        solo = [round(t.item(), 4) for t in allms]
        del gs
        cache.close("solo")
    sampler = ClockSampler(ctx.local)   # NVML set-up takes milliseconds and differs between ranks: outside the bracket
    ctx.barrier()
    sampler.start()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    evs[0].record()
    for k in range(K):
        xs = decode_round(xs)
        if world > 1 and stage.handoff == "peer_store":
            xs = None
        pos += 1
        evs[k + 1].record()
    stage.drain()
    torch.cuda.synchronize()
    clocks = sampler.finish()
    if world > 1:
        dist.barrier()
    stage.check_errors()
    total_ms = ctx.max_over_ranks(evs[0].elapsed_time(evs[K]))
    lat = sorted(evs[k].elapsed_time(evs[k + 1]) for k in range(K))
    launches = stage.launches_per_round * K
    ms_per_step = total_ms / K
    tokens_per_step = batch * n_mb
    value = tokens_per_step / (ms_per_step * 1e-3)
    ctx_mid = prompt + W + K // 2
    sb = decode_step_bytes(shape, len(layer_ids), batch, ctx_mid, w8=w8)
    stage_gbs = sb["total"] / (ms_per_step / n_mb * 1e-3) / 1e9
    res = {
        "model": shape.name, "weights": ("int8 rows + fp32 scales, " + ("int8 activations per token, threshold %g (LLM.int8)" % args.a8_threshold if a8 else "bf16 activations")) if w8 else "bf16", "layers_total": n_total, "layers_per_stage": len(layer_ids), "batch": batch, "micro_batches": n_mb,
        "prompt": prompt, "value": round(value, 1), "ms_per_step": round(ms_per_step, 4),
        "ms_per_stage_step": round(ms_per_step / n_mb, 4), "p50_ms_per_token": round(lat[len(lat) // 2] / n_mb, 4),
        "tokens_per_step": tokens_per_step, "steps": K, "launches": int(launches), "clocks": clocks,
        "handoff": stage.handoff, "handoff_fallback_reason": stage.handoff_error,
        "launch": ("eager" if args.no_graph else "cuda_graph_per_round" if round_graph else "cuda_graph"),
        "round_graph_error": stage.round_error,
        "stage_roofline": {"bound": "hbm", "achieved": round(stage_gbs, 1), "peak": hbm_peak, "unit": "GB/s",
                           "frac": round(stage_gbs / hbm_peak, 4), "bytes_per_stage_step": sb["total"],
                           "note": "whole decode stage-step (all kernels + gaps), SURVEY 8d algorithmic bytes"},
        "prefill_first_call_ms": round(prefill_ms, 3),
        "prefill_first_call_tokens_per_sec": round(n_mb * batch * prompt / (prefill_ms * 1e-3), 1),
        "prefill": prefill, "chunk": chunk,
    }
    if solo is not None and world > 1:
        res["stage_step_alone_ms"] = solo
        res["pipeline_vs_slowest_stage"] = round(max(solo) / (ms_per_step / n_mb), 4)
        res["stage_step_alone_note"] = ("each stage's decode step replayed without hand-off (same layers, batch and context), all "
                                        "ranks at the same time, per rank; pipeline_vs_slowest_stage = max of these / the "
                                        "pipeline's time per stage-step")
    if world > 1:   # whole-pipeline prefill: all stages' FLOPs over the max-over-ranks time, against N x the peak
        fl = n_mb * prefill_flops(shape, n_total, batch, prompt)
        n_chunks = -(-prompt // chunk)
        res["prefill_pipeline"] = {"ms": round(prefill_ms, 3), "chunk_tokens": chunk, "chunks": n_chunks,
                                   "sequences": n_mb * batch, "tflops": round(fl / (prefill_ms * 1e-3) / 1e12, 1),
                                   "frac_of_n_gpu_peak": round(fl / (prefill_ms * 1e-3) / 1e12 / (tc_peak * world), 4),
                                   "fill_drain_bound": round(n_chunks * n_mb / (n_chunks * n_mb + world - 1), 4)}

    # ---- instrumented pass: per-kernel-class time with CUDA events inside dli_span_forward ----
    if full:
        saved, stage.rings = stage.rings, {}          # events cannot be recorded inside a replayed graph: eager rounds
        if xs is None and first:
            xs = [syn.make_hidden(shape, batch, 1, seed=7 + m, device=dev) for m in range(n_mb)]
        blk.profile(True)
        for _ in range(n_prof):
            xs = stage.run_round(gids, xs, batch, 1, pos, feed_back=True)
            pos += 1
        prof = blk.profile_read()
        blk.profile(False)
        stage.rings = saved
        gemm_ms = sum(prof[c][0] for c in GEMM_CLASSES)
        gemm_launches = sum(prof[c][1] for c in GEMM_CLASSES)
        gemm_bytes_per_launch = sum(sb["per_gemm_per_layer"].values()) / 4.0
        gemm_ms_per_launch = gemm_chain_time(blk, shape, batch, dev)
        achieved = gemm_bytes_per_launch / (gemm_ms_per_launch * 1e-3) / 1e9
        res["roofline"] = {
            "bound": "hbm", "kernel": "gemm_tcgen05_kernel (swap-AB split-K weight streaming; 4 launches/layer)",
            "achieved": round(achieved, 1), "peak": hbm_peak, "unit": "GB/s", "frac": round(achieved / hbm_peak, 4),
            "traffic": measured_traffic(shape.name, batch), "peak_source": f"MEASURED_PEAKS.json ({peak_kind}, hbm_gbs)",
            "bytes_per_launch": gemm_bytes_per_launch, "ms_per_launch": round(gemm_ms_per_launch, 5),
            "how": "CUDA events around the span's own 4 GEMMs/layer x all layers launched back to back through "
                   "dli_gemm (programmatic dependent launch active, as in the step)",
            "ms_per_launch_serialised": round(gemm_ms / max(1, gemm_launches), 5),
            "share_of_step": round(gemm_ms / max(1e-9, sum(v[0] for v in prof.values())), 4)}
        res["kernel_ms_per_step"] = {c: round(v[0] / (n_prof * n_mb), 4) for c, v in prof.items()}
        res["kernel_ms_per_step_sum"] = round(sum(v[0] for v in prof.values()) / (n_prof * n_mb), 4)

    # ---- e2e: the sequential client closed over the stages (token ring), tokens fed from the host every step ----
    if want_e2e:
        res["e2e"] = e2e_token_ring(ctx, stage, shape, gids, batch, total_ctx, n_e2e, len(layer_ids))
        if world == 1 and graph is not None:
            res["e2e_span_api"] = e2e_span_api(graph, shape, batch, dev, n_e2e)

    # ---- configs[1] also names batch 1: per-token latency of the same span with ONE sequence (one GPU only) ----
    if world == 1 and full and graph is not None and batch != 1:
        x1 = syn.make_hidden(shape, 1, prompt, seed=123, device=dev)
        blk("b1", x1, past_key_value=cache)
        n1 = min(K, 32)
        g1 = blk.graphed_decode("b1", cache, prompt + n1 + 8)
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
        sb1 = decode_step_bytes(shape, len(layer_ids), 1, prompt + 3 + n1 // 2)
        res["decode_batch1"] = {"tokens_per_sec": round(1e3 / ms1, 1), "ms_per_step": round(ms1, 4),
                                "p50_ms_per_token": round(lat1[n1 // 2], 4), "steps": n1,
                                "stage_roofline_frac": round(sb1["total"] / (ms1 * 1e-3) / 1e9 / hbm_peak, 4),
                                "bytes_per_step": sb1["total"]}
        cache.close("b1")
    ctx.barrier()
    del stage, blk, cache, graph
    free_cuda()
    return res


def e2e_span_api(graph, shape, batch, dev, n_e2e):
    """One GPU, the span API alone: LlamaBlock.graphed_decode(...).step with the hidden states copied from pinned host
    memory and the output copied back, synchronised every step (what round 1 reported as e2e)."""
    hin = [torch.empty(batch, 1, shape.hidden_size, dtype=torch.bfloat16).pin_memory() for _ in range(2)]
    hout = torch.empty(batch, 1, shape.hidden_size, dtype=torch.bfloat16).pin_memory()
    for t in hin:
        t.copy_(syn.make_hidden(shape, batch, 1, seed=3))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(n_e2e):
        y = graph.step(hin[k & 1].to(dev, non_blocking=True))
        hout.copy_(y, non_blocking=True)
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    nbytes = batch * shape.hidden_size * 2
    return {"value": round(batch * n_e2e / dt, 1), "unit": UNIT, "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes,
            "steps": n_e2e, "how": "LlamaBlock.graphed_decode(...).step per step, hidden states H2D from pinned memory, "
                                   "output D2H, synchronised every step (wall clock); no embedding, no head"}


def e2e_token_ring(ctx: Ctx, stage, shape, gids, batch, total_ctx, n_e2e, layers_per_stage):
    """The client's loop with host buffers.  Stage 0 embeds device-resident token ids in its entry kernel, the last
    stage ends in final norm -> lm_head -> argmax and stores the ids into stage 0's inbox.  Every step and micro-batch,
    the host on stage 0 (1) waits for that micro-batch's ids of the previous step (side stream: flag wait + D2H into
    pinned memory + event), (2) copies them back H2D as the next input, (3) launches the micro-batch's stage step --
    while the other micro-batches are in flight on the other stages.  Wall clock over n_e2e steps, max over ranks."""
    rank, world, dev, dist = ctx.rank, ctx.world, ctx.dev, ctx.dist
    n_mb = len(gids)
    mode = stage.enable_graphs(gids, total_ctx, batch=batch, ring="tokens")
    first = rank == 0
    if first:
        side = torch.cuda.Stream(device=dev, priority=-1)
        pinned = [torch.zeros(batch, dtype=torch.int32).pin_memory() for _ in range(n_mb)]
        for m, t in enumerate(pinned):
            t.copy_(syn.make_prompt(shape, batch, 1, seed=500 + m)[:, 0].to(torch.int32))
        done = [torch.cuda.Event() for _ in range(n_mb)]
    ring = stage.rings["tokens"]

    def one_step(timed: bool):
        for m in range(n_mb):
            if first:
                if ring.steps > 0:
                    done[m].synchronize()                              # (1) ids of the previous step are on the host
                ring.tok_inbox[m].copy_(pinned[m], non_blocking=True)  # (2) H2D: the next input
            stage.step_one("tokens", m)                                # (3)
        stage.end_round("tokens")
        if first:
            with torch.cuda.stream(side):
                for m in range(n_mb):
                    stage.wait_result("tokens", m, side)
                    pinned[m].copy_(ring.tok_inbox[m], non_blocking=True)
                    done[m].record(side)

    for _ in range(3):
        one_step(False)
    if first:
        for ev in done:
            ev.synchronize()
    ctx.barrier()
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        one_step(True)
    if first:
        for ev in done:
            ev.synchronize()
    stage.drain()
    dt = ctx.max_over_ranks(time.perf_counter() - t0)
    stage.check_errors()
    sb = decode_step_bytes(shape, layers_per_stage, batch, 0, head=True)
    return {"value": round(n_mb * batch * n_e2e / dt, 1), "unit": UNIT, "h2d_bytes_per_step": n_mb * batch * 4,
            "d2h_bytes_per_step": n_mb * batch * 4, "steps": n_e2e, "handoff": mode,
            "last_stage_head_bytes_per_step": sb["head"],
            "how": f"sequential client over {world} stage(s), {n_mb} micro-batch(es) in flight: token ids D2H to pinned host "
                   "memory and H2D back every step and micro-batch (the host waits only for the micro-batch it is about to "
                   "feed); embedding in stage 0's entry kernel, final norm + lm_head + argmax as the last stage's last three "
                   "kernels; wall clock, max over ranks"}


# ---------------------------------------------------------------------------------------------------------
# pipeline parity: the N-stage data path against one span and the oracle, before any timing
# ---------------------------------------------------------------------------------------------------------
def pipeline_parity(ctx: Ctx):
    from distributed_llm_inference_b200.client import SequentialClient
    from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
    from distributed_llm_inference_b200.parallel.pipeline import PipelineStage

    args, rank, world, dev, dist = ctx.args, ctx.rank, ctx.world, ctx.dev, ctx.dist
    shape = syn.SHAPES[MODEL]
    n_total = args.layers or shape.num_hidden_layers
    layer_ids = stage_layers(n_total, rank, world)
    first, last = rank == 0, rank == world - 1
    B, T, n_mb, n_dec = 4, 128, world, 4
    blk = build_span(shape, layer_ids, dev)
    cache = PartialLlamaSinkCache(1 << 20, 4, capacity_tokens=4 * n_mb * B * (T + 64) + 4096)
    cw = syn.make_decidable_client_weights(shape) if (first or last) else {}
    client_w = {}
    if first:
        client_w["embed_tokens"] = cw["embed_tokens"].to(dev)
    if last:
        client_w.update(norm=cw["norm"].to(dev), lm_head=cw["lm_head"].to(dev), norm_eps=1e-6)
    stage = PipelineStage(blk, cache, rank, world, dev, client=client_w)
    rec = {"stages": world, "batch": B, "prompt": T, "micro_batches": n_mb, "decode_rounds": n_dec}

    def run_ring(kind, gids, first_inputs):
        """prefill through the stages (NCCL), then n_dec decode rounds over ring `kind`; -> per round, per micro-batch
        what came back to stage 0 (clones), and the prefill outputs."""
        if kind == "tokens":
            pf_in = [client_w["embed_tokens"][p] for p in first_inputs["prompt_ids"]] if first else None
        else:
            pf_in = first_inputs["prompt_hidden"] if first else None
        pf_out = stage.run_round(gids, pf_in, B, T, 0, feed_back=True)
        mode = stage.enable_graphs(gids, T + n_dec + 8, batch=B, peer_store=not args.nccl_handoff, ring=kind)
        outs, xs = [], (first_inputs["step0"] if first else None)
        for r in range(n_dec):
            if r == 2 and mode == "peer_store":
                stage.capture_round(gids, ring=kind)
            got = stage.run_round(gids, xs, B, 1, T + r, feed_back=True, ring=kind)
            stage.drain()
            if first:
                outs.append([t.clone() for t in got])
                xs = None if mode == "peer_store" else got
            else:
                xs = None
        stage.check_errors()
        return mode, pf_out, outs

    # --- hidden-state ring ---
    gids_h = [f"par_h{m}" for m in range(n_mb)]
    ins_h = {"prompt_hidden": [syn.make_hidden(shape, B, T, seed=300 + m, device=dev) for m in range(n_mb)],
             "step0": [syn.make_hidden(shape, B, 1, seed=400 + m, device=dev) for m in range(n_mb)]}
    mode, pf_out, outs_h = run_ring("hidden", gids_h, ins_h)
    rec["handoff"] = mode
    # --- token ring (needs peer memory) ---
    outs_t = None
    if mode == "peer_store":
        gids_t = [f"par_t{m}" for m in range(n_mb)]
        ins_t = {"prompt_ids": [syn.make_prompt(shape, B, T, seed=600 + m, device=dev) for m in range(n_mb)],
                 "step0": [syn.make_prompt(shape, B, 1, seed=700 + m, device=dev)[:, 0].to(torch.int32) for m in range(n_mb)]}
        _, _, outs_t = run_ring("tokens", gids_t, ins_t)
    ctx.barrier()
    del stage, blk, cache
    free_cuda()

    ok = True
    if first:   # the same layers as ONE span on this GPU, teacher-forced with what the pipeline fed its stage 0
        full = build_span(shape, list(range(n_total)), dev)
        fcache = PartialLlamaSinkCache(1 << 20, 4, capacity_tokens=4 * n_mb * B * (T + 64) + 4096)

        def cmp(a, b):
            a, b = a.float(), b.float()
            return (a - b).abs().max().item(), b.abs().max().item()

        worst, scale_at, pf_worst = 0.0, 1.0, 0.0
        for m in range(n_mb):
            res = full(gids_h[m], ins_h["prompt_hidden"][m], past_key_value=fcache, output_hidden_states=(m == 0))
            if m == 0:
                after2 = res[1][2].float().cpu()
            e, s = cmp(pf_out[m], res[0])
            pf_worst = max(pf_worst, e / s)
            x = ins_h["step0"][m]
            for r in range(n_dec):
                y = full(gids_h[m], x, past_key_value=fcache, cache_position=torch.tensor([T + r], device=dev))[0]
                e, s = cmp(outs_h[r][m], y)
                if e / s > worst:
                    worst, scale_at = e / s, s
                x = outs_h[r][m]                  # teacher-forced: the pipeline's own next input
        rec["hidden"] = {"max_abs_err_over_scale": round(worst, 5), "scale": round(scale_at, 4),
                         "prefill_max_abs_err_over_scale": round(pf_worst, 5), "tolerance": 3e-2,
                         "ok": bool(worst <= 3e-2 and pf_worst <= 3e-2),
                         "what": f"{world}-stage pipeline output vs the same {n_total} layers as one span on rank 0, prefill and "
                                 f"{n_dec} decode rounds (rounds 2.. replayed from the round graph), max-abs error / max-abs"}
        ok &= rec["hidden"]["ok"]
        # the one-span reference itself against the CPU oracle over its first two layers
        from oracle import span_oracle as so

        sd = full.state_dict()                  # the very values the GPU run used (generated on the device), up-cast
        Wc = [{k: sd[f"layers.{j}.{k}.weight"].float().cpu() for k in syn.LINEARS + syn.NORMS} for j in (0, 1)]
        oshape = so.SpanShape(shape.hidden_size, shape.intermediate_size, shape.num_attention_heads,
                              shape.num_key_value_heads, shape.head_dim, shape.rope_theta)
        osp = so.OracleSpan(oshape, [so.LayerWeights(*(w[k] for k in syn.LINEARS + syn.NORMS)) for w in Wc], "eager", "float")
        with torch.no_grad():
            ref2 = osp.forward("o", ins_h["prompt_hidden"][0].float().cpu())[0]
        e2 = (after2 - ref2).abs().max().item() / ref2.abs().max().item()
        rec["oracle_2_layers"] = {"max_abs_err_over_scale": round(e2, 5), "tolerance": 3e-2, "ok": bool(e2 <= 3e-2),
                                  "what": "hidden states after layer 2 of the one-span run vs the fp32 CPU oracle, prefill"}
        ok &= rec["oracle_2_layers"]["ok"]
        if outs_t is not None:
            cwd = {k: v.to(dev) for k, v in cw.items()}
            scl = SequentialClient(cwd["embed_tokens"], cwd["norm"], cwd["lm_head"], [full], [fcache])
            n_same = n_all = n_dec_ok = n_decidable = 0
            min_margin = float("inf")
            for m in range(n_mb):
                scl.step(gids_t[m], ins_t["prompt_ids"][m], 0)
                feed = ins_t["step0"][m].long()
                for r in range(n_dec):
                    logits = scl.step(gids_t[m], feed[:, None], T + r)
                    top2 = logits.topk(2, dim=-1).values
                    margin = top2[:, 0] - top2[:, 1]
                    want = scl.last_tokens.long()
                    got = outs_t[r][m].long()
                    same = want == got
                    decidable = margin > 0.25          # well above the 2e-2 logit tolerance of a bf16 run
                    n_same += int(same.sum().item())
                    n_all += want.numel()
                    n_decidable += int(decidable.sum().item())
                    n_dec_ok += int((same & decidable).sum().item())
                    min_margin = min(min_margin, margin.min().item())
                    feed = got
            rec["tokens"] = {"identical": n_same, "of": n_all, "decidable": n_decidable, "identical_where_decidable": n_dec_ok,
                             "min_top2_margin": round(min_margin, 4),
                             "ok": bool(n_dec_ok == n_decidable and n_decidable >= n_all // 2),
                             "what": f"token ids from the {world}-stage token ring (embedding -> stages -> head -> argmax) vs "
                                     "the sequential client over one span, teacher-forced; identity is required wherever the "
                                     "one-span run's top-1/top-2 logit margin exceeds 0.25"}
            ok &= rec["tokens"]["ok"]
        del full, fcache
        free_cuda()
    ok = ctx.all_ok(ok)
    rec["ok"] = ok
    return rec


def run_ours(args):
    ctx = Ctx(args)
    rank, world = ctx.rank, ctx.world
    K, W = args.steps, max(args.warmup, 3)
    if args.exp_c5_chunks:   # experiments: config C5 alone, once per chunk size
        out = {}
        for c in (int(v) for v in args.exp_c5_chunks.split(",")):
            r = run_workload(ctx, "70b", 1, 8192, 1, None, 4, 3, prefill_chunk=c)
            out[str(c)] = {k: r.get(k) for k in ("prefill_pipeline", "value", "ms_per_step")}
        if rank == 0:
            print(json.dumps({"c5_chunks": out, "n_gpus": world}), flush=True)
        if world > 1:
            ctx.dist.destroy_process_group()
        return
    parity = None
    if world > 1 and not args.no_parity:
        parity = pipeline_parity(ctx)
    n_mb = args.micro_batches or world   # micro-batches in flight (default: the pipeline depth)
    main = run_workload(ctx, MODEL, BATCH, PROMPT, n_mb, args.layers, K, W, prefill_chunk=args.prefill_chunk, full=not (args.w8 or args.a8),
                        want_e2e=not args.no_e2e, page_tokens=args.page_tokens, w8=args.w8 or args.a8, a8=args.a8)
    w8_rec = a8_rec = None
    if world == 1 and not (args.w8 or args.a8) and not args.no_w8:
        # the same workload with the weights stored as row-wise int8 (the reference's optimised mode keeps its weights
        # in that format, utils/model.py:93-123): a second, clearly labelled record -- the headline stays bf16
        r8 = run_workload(ctx, MODEL, BATCH, PROMPT, n_mb, args.layers, min(K, 32), 3, w8=True)
        w8_rec = {k: r8[k] for k in ("weights", "value", "ms_per_step", "p50_ms_per_token", "stage_roofline", "launches",
                                     "prefill_first_call_tokens_per_sec")}
        w8_rec["vs_bf16"] = round(r8["value"] / main["value"], 3)
        # ... and with the activations quantised per token as well: the arithmetic of that mode itself (LLM.int8,
        # convert_to_optimized_block(quantize=True)); int8 x int8 on the tensor cores, no widening of the weights
        ra = run_workload(ctx, MODEL, BATCH, PROMPT, n_mb, args.layers, min(K, 32), 3, w8=True, a8=True)
        a8_rec = {k: ra[k] for k in ("weights", "value", "ms_per_step", "p50_ms_per_token", "stage_roofline", "launches",
                                     "prefill_first_call_tokens_per_sec")}
        a8_rec["vs_bf16"] = round(ra["value"] / main["value"], 3)
    configs = None
    if world > 1 and not args.no_configs and MODEL == "7b" and not args.layers:
        configs = {}
        Kc, Wc = min(K, 16), 3
        if world == 2:
            configs["c3_13b_2x20_b8"] = run_workload(ctx, "13b", 8, 512, 2, None, Kc, Wc)
        if world == 8:
            configs["c4_i_70b_8x10_1x32"] = run_workload(ctx, "70b", 32, 512, 1, None, Kc, Wc)
            configs["c4_ii_70b_8x10_8x4"] = run_workload(ctx, "70b", 4, 512, 8, None, Kc, Wc)
            configs["c4_iii_70b_8x10_8x32"] = run_workload(ctx, "70b", 32, 512, 8, None, Kc, Wc, want_e2e=True)
            # the same schedule in the reference's int8 mode (LLM.int8: int8 weights, per-token int8 activations)
            configs["c4_iii_70b_8x10_8x32_llm_int8"] = run_workload(ctx, "70b", 32, 512, 8, None, Kc, Wc, w8=True, a8=True)
        if world in (4, 8):
            configs[f"c5_70b_prefill_8192_{world}_stages"] = run_workload(ctx, "70b", 1, 8192, 1, None, 4, 3, prefill_chunk=PREFILL_CHUNK_TOKENS)
    if rank == 0:
        shape = syn.SHAPES[MODEL]
        cpu = (cpu_reference_decode(shape, main["layers_total"], BATCH * n_mb, PROMPT, budget_s=args.cpu_budget)
               if args.cpu_budget > 0 and world == 1 else None)
        e2e = main.get("e2e")
        if e2e is not None:
            e2e = dict(e2e)
            e2e["frac_of_value"] = round(e2e["value"] / main["value"], 4)
        line = {
            "metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": main["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"{shape.name}-shaped span ({main['layers_total']} layers) KV-cached decode, batch {BATCH} x "
                                   f"{n_mb} micro-batch(es) in flight, {PROMPT}-token prompt, context "
                                   f"{PROMPT + W}..{PROMPT + W + K}; random-init weights, synthetic hidden states",
                       "global_batch": main["tokens_per_step"], "seq_len": PROMPT, "parallelism": f"pp{world}",
                       "l2": "inputs larger than L2 (12.9 GB of weights streamed per step)", "page_tokens": args.page_tokens,
                       "launch": main["launch"], "round_graph_error": main["round_graph_error"],
                       "handoff": main["handoff"], "handoff_fallback_reason": main["handoff_fallback_reason"]},
            "p50_ms_per_token": main["p50_ms_per_token"],
            "prefill_tokens_per_sec": (main["prefill"]["tokens_per_sec"] if main.get("prefill") else
                                       main["prefill_first_call_tokens_per_sec"]),
            "prefill_first_call_tokens_per_sec": main["prefill_first_call_tokens_per_sec"],
            "prefill_note": ("pipelined over the stages in chunks of %d tokens, warm" % main["chunk"]) if world > 1 else
                            "warm chunk (details in `prefill`); the first call also pays one-time descriptor/workspace set-up",
            "prefill_pipeline": main.get("prefill_pipeline"), "prefill": main.get("prefill"),
            "gpu_launches": main["launches"],
            "clocks": main["clocks"],
            "roofline": main.get("roofline"),
            "stage_roofline": main["stage_roofline"],
            "kernel_ms_per_step": main.get("kernel_ms_per_step"), "kernel_ms_per_step_sum": main.get("kernel_ms_per_step_sum"),
            "e2e": e2e, "e2e_span_api": main.get("e2e_span_api"),
            "cpu_baseline": cpu, "decode_batch1": main.get("decode_batch1"),
            "stage_step_alone_ms": main.get("stage_step_alone_ms"),
            "pipeline_vs_slowest_stage": main.get("pipeline_vs_slowest_stage"),
            "decode_int8_weights": w8_rec, "decode_llm_int8": a8_rec,
            "pipeline_parity": parity, "configs": configs,
        }
        print(json.dumps(line), flush=True)
    bad = parity is not None and not parity["ok"]
    if world > 1:
        ctx.dist.destroy_process_group()
    if bad:
        sys.exit(3)


def run_reference(args):
    """The reference arm: the reference's OWN CPU implementation of the path on the host cores -- its LlamaBlock /
    decoder layer / sink cache files, copied at build time from /root/reference to the git-ignored oracle/_ref and
    imported through oracle/ref_shim.py (two import shims for the newer transformers, three crash repairs; no source
    edited) -- on this arm's config, metric and unit.  Without oracle/_ref the oracle port is timed and the record
    says so.  Two measurements: BASELINE configs[0] in full (tiny model, 2 spans, prefill + 32 greedy steps), and for
    the headline one real layer of the 7B shape (real 512-token prefill, then decode steps) x 32 layers, EXTRAPOLATED
    because fp32 weights of the full model and minutes per token do not fit a bounded run.  --gpus N scales the global
    batch to the GPU arm's (16 x N sequences)."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    shape = syn.SHAPES[MODEL]
    K, W = args.steps, args.warmup
    gbatch = BATCH * max(1, args.gpus)
    budget = min(45.0, max(5.0, 3.0 * (K + W)))
    c1 = cpu_reference_config1()
    cpu = cpu_reference_decode(shape, shape.num_hidden_layers, gbatch, PROMPT, budget_s=budget, real_prefill=True)
    line = {"impl": "reference", "metric": METRIC, "value": round(cpu["value"], 3), "unit": UNIT, "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": round(gbatch / cpu["value"] * 1e3, 2), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"Llama-2-7B-shaped span (32 layers) KV-cached decode, batch {gbatch}, "
                                   f"{PROMPT}-token context, CPU ({cpu['kind']}; one layer timed, x32 extrapolated)",
                       "global_batch": gbatch, "seq_len": PROMPT, "parallelism": "cpu"},
            "cpu_baseline": cpu, "config1_full": c1,
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
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the pipeline_parity record")
    ap.add_argument("--no-configs", action="store_true", help="N > 1: skip the BASELINE configs 3-5 records")
    ap.add_argument("--no-e2e", action="store_true", help="skip the token-ring e2e measurement")
    ap.add_argument("--exp-c5-chunks", default=None, help="experiments: run config C5 alone with these chunk sizes (comma-separated)")
    ap.add_argument("--w8", action="store_true", help="run the main workload with row-wise int8 weights (experiments)")
    ap.add_argument("--a8", action="store_true", help="with int8 weights also quantise the activations per token (LLM.int8; experiments)")
    ap.add_argument("--a8-threshold", type=float, default=5.0, help="LLM.int8 outlier threshold of --a8 (0: one plane, no outliers)")
    ap.add_argument("--no-w8", action="store_true", help="one GPU: skip the decode_int8_weights record")
    ap.add_argument("--no-chain", action="store_true", help="experiments: a stream barrier in front of every stage step instead "
                                                            "of chaining the micro-batches' steps with programmatic dependent launch")
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
