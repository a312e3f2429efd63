"""``InferenceWorker`` — hosts one span of a model behind an ``InferenceBackend``.

Reference: distributed_llm_inference/server/worker.py:9-23 is a stub (``__init__`` body is commented pseudo-code,
``run`` is ``pass``); its signature ``(model, block_index_start, block_index_end, ...)`` and the comment
``self.blocks[...] = load_block()`` say what it was meant to do.  This version does exactly that: load the span
``[block_index_start, block_index_end)`` from a local checkpoint directory, put it on the GPU, give it a paged
sink cache and wrap it in the backend whose pool the server loop (server/server.py) calls.
"""
from __future__ import annotations

from typing import Optional, TypedDict

import torch

from distributed_llm_inference_b200.models.llama.cache import PartialLlamaSinkCache
from distributed_llm_inference_b200.server.backend import BatchTensorDescriptor, InferenceBackend
from distributed_llm_inference_b200.utils.model import convert_to_optimized_block, load_block


class Block(TypedDict):          # server/worker.py:4-6
    block_index: int
    block_id: str


class InferenceWorker:
    def __init__(self, model: str, block_index_start: int, block_index_end: int, *, window_length: int = 4096,
                 num_sink_tokens: int = 4, capacity_tokens: Optional[int] = None, rope: str = "literal",
                 attn_implementation: Optional[str] = None, device: Optional[str] = None):
        self.layer_ids = list(range(block_index_start, block_index_end))
        self.blocks = [Block(block_index=i, block_id=f"{model}.layers.{i}") for i in self.layer_ids]
        block = load_block(model, self.layer_ids, rope=rope, attn_implementation=attn_implementation)
        self.block = convert_to_optimized_block(block) if device is None else block.to(device)
        self.cache = PartialLlamaSinkCache(window_length, num_sink_tokens, capacity_tokens=capacity_tokens)
        h = self.block.config.hidden_size
        self.backend = InferenceBackend(
            f"span.{block_index_start}:{block_index_end}", self.block,
            args_schema=(BatchTensorDescriptor((1, h), torch.bfloat16),), cache=self.cache)

    def forward(self, generation_id: str, hidden_states: torch.Tensor, cache_position: Optional[torch.Tensor] = None):
        return self.backend.forward(hidden_states, generation_id=generation_id, cache_position=cache_position)[0]

    def run(self):
        """-> the backend's inference pool; ``Server.run`` (server/server.py) takes tasks from its queue and submits them
        here one at a time.  Transport between machines (hivemind in the reference's imports) is outside this repository."""
        return self.backend.get_pools()
