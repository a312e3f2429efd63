"""world_size-2 gloo test of the pipeline hand-off logic on CPU (fake spans; the CUDA span has no CPU path)."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class FakeSpan:
    """y = x * scale + pos + per-generation call count (so order, positions and sessions are all observable)."""

    def __init__(self, scale):
        self.scale, self.calls, self.last_launch_count = scale, {}, 7

    def __call__(self, gid, x, past_key_value=None, cache_position=None):
        n = self.calls.get(gid, 0)
        self.calls[gid] = n + 1
        return (x * self.scale + float(cache_position[0]) + n,)


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    from distributed_llm_inference_b200.parallel.pipeline import PipelineStage

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    stage = PipelineStage(FakeSpan(2.0 + rank), None, rank, world, "cpu", hidden_size=4, dtype=torch.float32)
    gids = ["a", "b"]
    xs = [torch.full((3, 1, 4), 1.0), torch.full((3, 1, 4), 10.0)] if rank == 0 else None
    hist = []
    for step in range(3):
        xs = stage.run_round(gids, xs, 3, 1, 100 + step, feed_back=True)
        if step == 0:       # no peer-store hand-off on this backend: every stage must agree NOT to use a round graph
            assert stage.capture_round(gids) is False and stage._round is None
        if rank == 0:
            hist.append([float(t[0, 0, 0]) for t in xs])
    final = stage.run_round(gids, xs, 3, 1, 103, feed_back=False)
    launches = stage.launches_per_round
    # chunked prefill: a 10-token prompt in chunks of 4 -> chunks at positions 0, 4, 8 (the last one ragged, 2 tokens)
    prompts = [torch.arange(10, dtype=torch.float32).view(1, 10, 1).expand(3, 10, 4).contiguous()] if rank == 0 else None
    pre = stage.prefill(["p"], prompts, 3, 10, chunk_tokens=4)
    if rank == world - 1:
        out.put(("last", [float(t[0, 0, 0]) for t in final], launches))
        out.put(("prefill", [list(pre[0].shape), [float(v) for v in pre[0][0, :, 0]]], 0))
    if rank == 0:
        out.put(("first", hist, launches))
    stage._reap()
    dist.barrier()
    dist.destroy_process_group()


def test_two_stage_pipeline_over_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    got = dict()
    for _ in range(3):
        tag, vals, launches = q.get(timeout=120)
        got[tag] = (vals, launches)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)

    # replay the same schedule serially: stage0 (scale 2) then stage1 (scale 3), ring-closed
    def serial():
        xs, calls0, calls1, hist = [1.0, 10.0], [0, 0], [0, 0], []
        for step in range(4):
            ys = []
            for m in range(2):
                h = xs[m] * 2.0 + (100 + step) + calls0[m]; calls0[m] += 1
                h = h * 3.0 + (100 + step) + calls1[m]; calls1[m] += 1
                ys.append(h)
            xs = ys
            hist.append(list(ys))
        return hist
    ref = serial()
    assert got["first"][0] == ref[:3]
    assert got["last"][0] == ref[3]
    assert got["first"][1] == 14 and got["last"][1] == 14      # 2 micro-batches x 7 launches per stage
    # prefill: the last stage holds the LAST chunk (tokens 8, 9 at position 8; third call of generation "p" on each stage)
    shape, vals = got["prefill"][0]
    assert shape == [3, 2, 4]
    assert vals == [((t * 2.0 + 8 + 2) * 3.0 + 8 + 2) for t in (8.0, 9.0)]
