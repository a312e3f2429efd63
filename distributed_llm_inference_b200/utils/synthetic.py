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


# ---------------------------------------------------------------------------------------------------------
# "Decidable" client: greedy token identity between a bf16 and an fp32 run is only a meaningful assertion when
# the model's top-1/top-2 margins are well above the arithmetic error.  A model whose every tensor is
# N(0, 0.02^2) has margins down to 1e-3 (the reference's own bf16 run disagrees with its fp32
# run).  This recipe — fixed here, before any GPU result was looked at — keeps the decoder layers exactly as
# above and only shapes the client tensors the way a trained model is shaped: unit-scale embeddings, and an
# output head tied to the embedding through a fixed permutation of the vocabulary,
#     lm_head[perm[t]] = (alpha / sqrt(h)) * embed[t],
# so the hidden state of token t (its embedding plus everything the layers add) scores highest for perm[t].
# Logits of the other tokens stay O(alpha); the layers still move every logit, so the 2e-2 logit bound keeps
# testing the arithmetic, while the argmax is decided by a margin of order alpha * sqrt(h) / rms(residual).
# ---------------------------------------------------------------------------------------------------------
DECIDABLE_ALPHA = 0.5
DECIDABLE_SEED = 2468


def make_decidable_client_weights(shape: LlamaShape, device="cpu", dtype=torch.bfloat16, seed: int = DECIDABLE_SEED,
                                  alpha: float = DECIDABLE_ALPHA):
    gen = torch.Generator(device="cpu")
    gen.manual_seed(seed)
    emb = torch.empty(shape.vocab_size, shape.hidden_size, dtype=torch.float32).normal_(0, 1.0, generator=gen)
    emb = emb.to(torch.bfloat16)
    perm = torch.randperm(shape.vocab_size, generator=gen)
    head = torch.empty_like(emb)
    head[perm] = (emb.float() * (alpha / shape.hidden_size ** 0.5)).to(torch.bfloat16)
    norm = torch.ones(shape.hidden_size, dtype=torch.bfloat16)
    return {"embed_tokens": emb.to(dtype).to(device), "norm": norm.to(dtype).to(device), "lm_head": head.to(dtype).to(device)}


def write_checkpoint(model_dir: str, shape: LlamaShape, layer_weights: Dict[int, Dict[str, torch.Tensor]],
                     client: Optional[Dict[str, torch.Tensor]] = None, shard_layers: int = 0,
                     attn_implementation: str = "eager", max_position_embeddings: int = 8192) -> None:
    """A local Hugging-Face-style checkpoint directory (config.json + safetensors [+ index]) holding the given layers
    under the names the reference's loader looks for, ``model.layers.{global_idx}.<param>.weight``
    (distributed_llm_inference/utils/model.py:40,64-70), and the client tensors under ``model.embed_tokens.weight``,
    ``model.norm.weight``, ``lm_head.weight``.  ``shard_layers > 0`` splits the layers over several files behind a
    ``model.safetensors.index.json`` weight map."""
    import json
    import os

    from safetensors.torch import save_file

    os.makedirs(model_dir, exist_ok=True)
    cfg = shape.to_hf_config(attn_implementation, max_position_embeddings)
    cfg.architectures = ["LlamaForCausalLM"]
    cfg.save_pretrained(model_dir)
    tensors = {}
    for g, w in layer_weights.items():
        for name, t in w.items():
            tensors[f"model.layers.{g}.{name}.weight"] = t.contiguous().cpu()
    extra = {}
    if client is not None:
        extra = {"model.embed_tokens.weight": client["embed_tokens"].contiguous().cpu(),
                 "model.norm.weight": client["norm"].contiguous().cpu(),
                 "lm_head.weight": client["lm_head"].contiguous().cpu()}
    if shard_layers <= 0:
        save_file({**tensors, **extra}, os.path.join(model_dir, "model.safetensors"))
        return
    ids = sorted(layer_weights)
    groups = [ids[i:i + shard_layers] for i in range(0, len(ids), shard_layers)]
    n_files = len(groups) + (1 if extra else 0)
    weight_map = {}
    for fi, grp in enumerate(groups):
        fname = f"model-{fi + 1:05d}-of-{n_files:05d}.safetensors"
        part = {k: v for k, v in tensors.items() if int(k.split(".")[2]) in grp}
        save_file(part, os.path.join(model_dir, fname))
        weight_map.update({k: fname for k in part})
    if extra:
        fname = f"model-{n_files:05d}-of-{n_files:05d}.safetensors"
        save_file(extra, os.path.join(model_dir, fname))
        weight_map.update({k: fname for k in extra})
    with open(os.path.join(model_dir, "model.safetensors.index.json"), "w") as f:
        json.dump({"metadata": {}, "weight_map": weight_map}, f)


def read_client_weights(model_dir: str, device="cpu", dtype=torch.bfloat16) -> Dict[str, torch.Tensor]:
    """The client tensors of a checkpoint written by ``write_checkpoint`` (embedding, final norm, head)."""
    import json
    import os

    import safetensors

    names = {"model.embed_tokens.weight": "embed_tokens", "model.norm.weight": "norm", "lm_head.weight": "lm_head"}
    index = os.path.join(model_dir, "model.safetensors.index.json")
    if os.path.exists(index):
        with open(index) as f:
            wm = json.load(f)["weight_map"]
        files = {k: os.path.join(model_dir, wm[k]) for k in names}
    else:
        files = {k: os.path.join(model_dir, "model.safetensors") for k in names}
    out = {}
    for k, short in names.items():
        with safetensors.safe_open(files[k], framework="pt", device="cpu") as f:
            out[short] = f.get_tensor(k).to(dtype).to(device)
    return out
