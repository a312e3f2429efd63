"""Synthetic workloads: the Llama shapes named by BASELINE.json and seeded random-init weights.

There is no network for checkpoints, so every parity and bench run uses random-init weights of the
public Llama-2 shapes.  Each ``nn.Linear`` weight is drawn N(0, 0.02^2) (HF
``initializer_range``) with ``torch.manual_seed(1234 + global_layer_idx)`` in fp32 and rounded ONCE to
bf16; the CPU oracle up-casts those exact values, the GPU path uses them as they are.  RMSNorm
weights are 1 (HF default init) unless ``norm_jitter`` asks for a perturbation (parity tests use it so
that a dropped norm weight cannot hide).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional

import torch

LINEARS = ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj", "self_attn.o_proj",
           "mlp.gate_proj", "mlp.up_proj", "mlp.down_proj")
NORMS = ("input_layernorm", "post_attention_layernorm")


@dataclass(frozen=True)
class LlamaShape:
    name: str
    hidden_size: int
    intermediate_size: int
    num_hidden_layers: int
    num_attention_heads: int
    num_key_value_heads: int
    head_dim: int
    vocab_size: int = 32000
    rope_theta: float = 10000.0

    def linear_shapes(self) -> Dict[str, tuple]:
        h, i = self.hidden_size, self.intermediate_size
        q, kv = self.num_attention_heads * self.head_dim, self.num_key_value_heads * self.head_dim
        return {"self_attn.q_proj": (q, h), "self_attn.k_proj": (kv, h), "self_attn.v_proj": (kv, h),
                "self_attn.o_proj": (h, q), "mlp.gate_proj": (i, h), "mlp.up_proj": (i, h),
                "mlp.down_proj": (h, i)}

    def layer_weight_bytes(self, elt: int = 2) -> int:
        return elt * (sum(a * b for a, b in self.linear_shapes().values()) + 2 * self.hidden_size)

    def to_hf_config(self, attn_implementation: str = "eager", max_position_embeddings: int = 8192):
        from transformers import LlamaConfig

        cfg = LlamaConfig(
            hidden_size=self.hidden_size, intermediate_size=self.intermediate_size,
            num_attention_heads=self.num_attention_heads, num_key_value_heads=self.num_key_value_heads,
            num_hidden_layers=self.num_hidden_layers, vocab_size=self.vocab_size,
            max_position_embeddings=max_position_embeddings,
            rope_parameters={"rope_type": "default", "rope_theta": self.rope_theta},
            attention_bias=False, mlp_bias=False, pretraining_tp=1, head_dim=self.head_dim,
        )
        cfg._attn_implementation = attn_implementation
        return cfg


SHAPES = {
    "tiny": LlamaShape("tiny", 768, 2048, 12, 12, 12, 64),
    "7b": LlamaShape("llama2-7b", 4096, 11008, 32, 32, 32, 128),
    "13b": LlamaShape("llama2-13b", 5120, 13824, 40, 40, 40, 128),
    "70b": LlamaShape("llama2-70b", 8192, 28672, 80, 64, 8, 128),
    # small test shapes (not in BASELINE.json): GQA + hd=128 and hd=64 variants that run in seconds on CPU
    "test_gqa": LlamaShape("test_gqa", 512, 1408, 4, 8, 2, 64),
    "test_hd128": LlamaShape("test_hd128", 512, 1024, 4, 4, 2, 128),
}


def make_layer_weights(shape: LlamaShape, global_layer_idx: int, device="cpu", dtype=torch.bfloat16,
                       norm_jitter: float = 0.0, std: float = 0.02) -> Dict[str, torch.Tensor]:
    """One layer's tensors keyed by the reference's parameter names (without ``.weight``)."""
    dev = torch.device(device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + global_layer_idx)
    out: Dict[str, torch.Tensor] = {}
    for name, shp in shape.linear_shapes().items():
        w = torch.empty(shp, dtype=torch.float32, device=dev).normal_(0.0, std, generator=gen)
        out[name] = w.to(torch.bfloat16).to(dtype)
    for name in NORMS:
        w = torch.ones(shape.hidden_size, dtype=torch.float32, device=dev)
        if norm_jitter:
            w = w + norm_jitter * torch.empty_like(w).normal_(0.0, 1.0, generator=gen)
        out[name] = w.to(torch.bfloat16).to(dtype)
    return out


def make_span_weights(shape: LlamaShape, layer_ids: List[int], **kw) -> List[Dict[str, torch.Tensor]]:
    return [make_layer_weights(shape, g, **kw) for g in layer_ids]


def make_client_weights(shape: LlamaShape, device="cpu", dtype=torch.bfloat16, seed: int = 4321):
    """Harness-client tensors (embedding, final norm, untied lm_head) — SURVEY section 8d."""
    dev = torch.device(device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed)
    emb = torch.empty(shape.vocab_size, shape.hidden_size, dtype=torch.float32, device=dev).normal_(0, 0.02, generator=gen)
    head = torch.empty(shape.vocab_size, shape.hidden_size, dtype=torch.float32, device=dev).normal_(0, 0.02, generator=gen)
    norm = torch.ones(shape.hidden_size, dtype=torch.float32, device=dev)
    cast = lambda t: t.to(torch.bfloat16).to(dtype)
    return {"embed_tokens": cast(emb), "norm": cast(norm), "lm_head": cast(head)}


def make_prompt(shape: LlamaShape, batch: int, length: int, seed: int = 99, device="cpu") -> torch.Tensor:
    gen = torch.Generator(device="cpu")
    gen.manual_seed(seed)
    return torch.randint(0, shape.vocab_size, (batch, length), generator=gen).to(device)


def make_hidden(shape: LlamaShape, batch: int, length: int, seed: int = 7, dtype=torch.bfloat16,
                device="cpu", scale: float = 1.0) -> torch.Tensor:
    """Span-input activations ``[B, L, h]`` (already rounded to bf16 so every implementation sees the
    same values)."""
    gen = torch.Generator(device="cpu")
    gen.manual_seed(seed)
    x = torch.empty(batch, length, shape.hidden_size, dtype=torch.float32).normal_(0, scale, generator=gen)
    return x.to(torch.bfloat16).to(dtype).to(device)
