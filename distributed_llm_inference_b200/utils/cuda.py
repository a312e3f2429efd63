"""CUDA-graph helper with the reference's name and calling convention.

Reference: distributed_llm_inference/utils/cuda.py:6-77 (``make_inference_graphed_callable``), a forward-only
variant of ``torch.cuda.make_graphed_callables`` that the reference uses to hide the launch latency of three tiny
per-layer ops at decode (modules.py:28-34,130-144).  The B200 span has no such per-op graphs — a whole decode
step of a stage is ~8 fused kernels per layer issued by one C call — so this helper captures ANY callable whose
work is CUDA launches on the current stream (e.g. a bound ``LlamaBlock`` decode step with static inputs) and
replays it with the same contract: side-stream warm-up (cuda.py:28-34), capture (cuda.py:37-38), replay with
``copy_`` into the static inputs when the caller passes different tensors (cuda.py:50-57).
"""
from __future__ import annotations

import torch
from torch.utils._pytree import tree_flatten, tree_unflatten


def make_inference_graphed_callable(callable, sample_args, num_warmup_iters: int = 3):
    assert not isinstance(callable, torch.nn.Module), "pass module.forward (bound method), as the reference does"
    flat_args, _ = tree_flatten(sample_args)
    assert all(isinstance(a, torch.Tensor) for a in flat_args), "sample_args must contain only tensors"
    static_inputs = tuple(flat_args)
    n_user = len(sample_args)

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s), torch.no_grad():
        for _ in range(num_warmup_iters):
            out = callable(*sample_args)
        del out
    torch.cuda.current_stream().wait_stream(s)

    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph), torch.no_grad():
        outputs = callable(*sample_args)
    flat_out, out_spec = tree_flatten(outputs)
    static_outputs = tuple(flat_out)

    def graphed(*user_args):
        flat, _ = tree_flatten(user_args)
        for i in range(n_user):
            if static_inputs[i].data_ptr() != flat[i].data_ptr():
                static_inputs[i].copy_(flat[i])
        graph.replay()
        return tree_unflatten([o.detach() for o in static_outputs], out_spec)

    graphed.graph = graph
    return graphed
