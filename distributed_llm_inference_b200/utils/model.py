"""Per-span checkpoint loading from a LOCAL model directory (no network), with the reference's function names.

Reference: distributed_llm_inference/utils/model.py:16-90 — ``load_block`` fetches, per layer, the safetensors shards
that hold ``model.layers.{global_idx}.*`` from the HF hub (``cached_file``; network), casts them to fp16 (:67-68) and
puts them on the CPU.  Here the same contract works offline: ``model_name`` is a directory that contains
``config.json`` and either ``model.safetensors.index.json`` + shards or a single ``model.safetensors``; only the
shards that hold the span's layers are opened; tensors are cast to bf16 (the span kernels' dtype; the reference's
fp16 cast is a dtype choice) and copied into the ``LlamaBlock``'s parameters, whose names are the
reference's (``layers.{j}.self_attn.q_proj.weight`` ...).  ``convert_to_optimized_block`` (bitsandbytes int8,
utils/model.py:93-123) moves the block to the GPU and optionally stores the weights as row-wise int8 (weight-only:
activations stay bf16).
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Optional

import torch

from distributed_llm_inference_b200.models.llama.model import LlamaBlock

INDEX_FILE_PATTERNS = ["model.safetensors.index.json", "model.safetensors"]   # utils/model.py:13 (safetensors only)


def get_sharded_block_state_from_file(file: str, block_prefix: str) -> Dict[str, torch.Tensor]:
    """utils/model.py:16-24."""
    import safetensors

    out = {}
    with safetensors.safe_open(file, framework="pt", device="cpu") as f:
        for key in f.keys():
            if key.startswith(block_prefix):
                out[key[len(block_prefix):]] = f.get_tensor(key)
    return out


def get_block_state_dict(model_dir: str, block_idx: int) -> Dict[str, torch.Tensor]:
    """utils/model.py:27-52 against a local directory."""
    prefix = f"model.layers.{block_idx}."
    index_path = os.path.join(model_dir, INDEX_FILE_PATTERNS[0])
    single = os.path.join(model_dir, INDEX_FILE_PATTERNS[1])
    if os.path.exists(index_path):
        with open(index_path) as f:
            index = json.load(f)
        if "weight_map" not in index:
            raise ValueError("Index file does not contain a weight map")          # utils/model.py:38-39
        files = sorted({v for k, v in index["weight_map"].items() if k.startswith(prefix)})
        paths = [os.path.join(model_dir, v) for v in files]
    elif os.path.exists(single):
        paths = [single]
    else:
        raise FileNotFoundError("Could not find index file for the provided repository")   # utils/model.py:34-35
    state: Dict[str, torch.Tensor] = {}
    for path in paths:
        state.update(get_sharded_block_state_from_file(path, prefix))
    return state


def _load_layer(model_dir: str, layer, layer_idx: int, dtype=torch.bfloat16):
    """utils/model.py:55-72."""
    state = get_block_state_dict(model_dir, layer_idx)
    with torch.no_grad():
        for name, param in layer.named_parameters():
            assert name in state, f"Parameter {name} not found in state_dict"     # utils/model.py:65
            value = state[name]
            if value.is_floating_point():
                value = value.to(dtype)
            if tuple(value.shape) != tuple(param.shape):
                raise ValueError(f"layer {layer_idx}: {name} has shape {tuple(value.shape)}, expected {tuple(param.shape)}")
            param.data.copy_(value)
    return layer


def load_block(model_name: str, layer_ids: List[int], use_quantized: bool = False, cache_dir: Optional[str] = None,
               token=False, *, rope: str = "literal", attn_implementation: Optional[str] = None) -> LlamaBlock:
    """utils/model.py:75-90: config -> ``LlamaBlock(config, layer_ids)`` -> per-layer weights (on the CPU; move the
    block to the GPU with ``.to("cuda")`` or ``convert_to_optimized_block``)."""
    if use_quantized:
        raise NotImplementedError("int8 (bitsandbytes) weights are outside the bf16 span kernels")
    from transformers import AutoConfig

    config = AutoConfig.from_pretrained(model_name, local_files_only=True)
    if attn_implementation is not None:
        config._attn_implementation = attn_implementation
    block = LlamaBlock(config, layer_ids=list(layer_ids), rope=rope)
    for layer in block.layers:
        _load_layer(model_name, layer, layer.layer_idx)
    return block


def convert_to_optimized_block(block: LlamaBlock, quantize: bool = False, threshold: float = 5.0,
                               activations: str = "int8") -> LlamaBlock:
    """utils/model.py:116-123: move the block to the GPU and, with ``quantize=True``, run its linear layers the way the
    reference's ``Linear8bitLt(threshold=threshold, has_fp16_weights=False)`` does (utils/model.py:102-108, "LLM.int8"):
    every weight stored as row-wise absmax int8 + one fp32 scale per output row, and at decode (chunks of <= 128 tokens)
    the ACTIVATIONS quantised per token row to int8 as well, the entries with ``|x| >= threshold`` kept out of the row's
    scale in a second int8 plane, the products int8 x int8 -> int32 on the tensor cores (``dli_gemm_i8``).  Longer
    chunks (prefill) multiply 16-bit activations by the dequantised weights.  ``activations="bf16"`` keeps 16-bit
    activations everywhere (weight-only int8, ``dli_gemm_w8``): closer to the unquantised model than the reference's
    mode, and slower at decode.  The reference converts regardless of ``quantize``; here ``quantize=False``
    keeps the bf16 span (the north star's dtype)."""
    if not torch.cuda.is_available():
        raise NotImplementedError("the B200 span needs a CUDA device" if not quantize else
                                  "Quantization is only supported on CUDA devices")      # utils/model.py:117-118
    block = block.to(torch.device("cuda"))
    if quantize:
        block.quantize_w8(activations=activations, threshold=threshold)
    return block
