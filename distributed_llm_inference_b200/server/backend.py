"""``InferenceBackend`` — the server-side wrapper around a span, usable without hivemind.

Reference: distributed_llm_inference/server/backend.py:11-51, a copy of hivemind's ``ModuleBackend.__init__``
(schemas, optional dummy forward to infer the output schema :31-35, one inference ``TaskPool`` :42, backward
disabled :44-48).  hivemind 1.1.10.post2 is not installable here, and the reference's version cannot carry the
``generation_id: str`` its own module requires (tensor-only schemas).  This mirror keeps the
constructor signature, ``forward(*flat_tensors)`` semantics (positional tensors then keyword tensors in schema
order, ``torch.no_grad``, flat tuple out), ``get_pools()`` and the ``NotImplementedError`` on backward, and carries
the generation id out-of-band of the tensor schema (``forward(..., generation_id=...)`` or ``session(...)``).
The output schema of a span is its input schema (``[B, L, h]`` in, ``[B, L, h]`` out), so no dummy forward on CPU
zeros is needed — the span has no CPU path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, Optional, Tuple, Union

import torch
from torch import nn

DUMMY_BATCH_SIZE = 3  # hivemind.utils.tensor_descr.DUMMY_BATCH_SIZE


@dataclass(frozen=True)
class BatchTensorDescriptor:
    """Minimal stand-in for hivemind's descriptor: per-sample shape (batch dim excluded) + dtype."""
    size: Tuple[int, ...]
    dtype: torch.dtype = torch.bfloat16
    device: Optional[Union[str, torch.device]] = None

    def make_zeros(self, batch_size: int) -> torch.Tensor:
        return torch.zeros(batch_size, *self.size, dtype=self.dtype, device=self.device)

    @classmethod
    def from_tensor(cls, t: torch.Tensor) -> "BatchTensorDescriptor":
        return cls(tuple(t.shape[1:]), t.dtype, t.device)


class TaskPool:
    """In-process pool: hivemind's forks a process and batches requests; calls into a span are serialised
    either way, so this one simply calls ``process_func``."""

    def __init__(self, process_func: Callable, name: str = "pool", **kwargs):
        self.process_func, self.name, self.kwargs = process_func, name, kwargs

    def submit_task(self, *args, **kwargs):
        return self.process_func(*args, **kwargs)


class InferenceBackend:
    def __init__(
        self,
        name: str,
        module: nn.Module,
        *,
        args_schema: Tuple[BatchTensorDescriptor, ...] = None,
        kwargs_schema: Dict[str, BatchTensorDescriptor] = None,
        outputs_schema: Union[BatchTensorDescriptor, Tuple[BatchTensorDescriptor, ...]] = None,
        cache=None,
        **kwargs,
    ):
        self.name, self.module, self.optimizer, self.scheduler = name, module, None, None
        self.args_schema = args_schema = tuple(args_schema or ())
        self.kwargs_schema = kwargs_schema = dict(kwargs_schema or {})
        assert args_schema or kwargs_schema, (
            "Module must take at least one positional or keyword input."
            " Did you forget to provide args_schema/kwargs_schema?"
        )
        if outputs_schema is None:
            # a span maps hidden states to hidden states of the same shape and dtype (model.py:76)
            outputs_schema = (args_schema[0] if args_schema else next(iter(kwargs_schema.values())),)
        self.forward_schema = (self.args_schema, self.kwargs_schema)
        self.outputs_schema = outputs_schema
        self.backward_schema = (self.forward_schema, self.outputs_schema)
        self.grad_inputs_schema = self.forward_schema
        self.cache = cache
        self._generation_id: Optional[str] = None
        self.inference_pool = TaskPool(self.forward, name=f"{self.name}_inference", **kwargs)

    def session(self, generation_id: str) -> "InferenceBackend":
        self._generation_id = generation_id
        return self

    def forward(self, *inputs: torch.Tensor, generation_id: Optional[str] = None,
                cache_position: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, ...]:
        gid = generation_id or self._generation_id
        n = len(self.args_schema)
        args, kw_vals = inputs[:n], inputs[n:]
        kwargs = dict(zip(self.kwargs_schema.keys(), kw_vals))
        if cache_position is not None:
            kwargs["cache_position"] = cache_position
        if self.cache is not None:
            kwargs.setdefault("past_key_value", self.cache)
        with torch.no_grad():
            out = self.module(gid, *args, **kwargs)
        return tuple(out) if isinstance(out, (tuple, list)) else (out,)

    def backward(self, *inputs: torch.Tensor):
        raise NotImplementedError("InferenceBackend does not support backward pass")

    def on_backward(self, batch_size: int) -> None:
        raise NotImplementedError("InferenceBackend does not support backward pass")

    def get_pools(self):
        return self.inference_pool
