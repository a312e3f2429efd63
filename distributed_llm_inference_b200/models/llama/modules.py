"""Parameter containers with the reference's module tree and state-dict names.

Reference: distributed_llm_inference/models/llama/modules.py:23-184.  There each module computes with
PyTorch ops; here the modules only OWN the weights, laid out the way the sm_100a kernels stream them:
``q_proj | k_proj | v_proj`` are row blocks of one ``[(H + 2 KV) hd, h]`` buffer and ``gate_proj | up_proj`` of
one ``[2 i, h]`` buffer (so one TMA descriptor feeds one fused GEMM), while ``*.weight`` stay addressable under
the reference's names — ``layers.{j}.self_attn.{q,k,v,o}_proj.weight``, ``layers.{j}.mlp.{gate,up,down}_proj
.weight``, ``layers.{j}.{input,post_attention}_layernorm.weight`` (modules.py:122-125 + HF LlamaAttention /
LlamaMLP) — as views into those buffers, so ``load_state_dict`` from a reference checkpoint fills them in place.
The arithmetic lives in ``LlamaBlock.forward`` -> ``dli_span_forward`` (one C call for the whole span).
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn


class _Weight(nn.Module):
    """Stand-in for a bias-free ``nn.Linear`` / ``LlamaRMSNorm``: just a ``weight`` parameter."""

    def __init__(self, *shape, dtype, ones: bool = False):
        super().__init__()
        init = torch.ones if ones else torch.empty
        self.weight = nn.Parameter(init(*shape, dtype=dtype), requires_grad=False)

    def forward(self, *a, **k):  # pragma: no cover - containers are never called
        raise RuntimeError("B200 span: layers are executed by LlamaBlock.forward (dli_span_forward), not one by one")


class OptimizedLlamaInferenceAttention(nn.Module):
    """Weights of reference ``OptimizedLlamaInferenceAttention`` (modules.py:23-114)."""

    def __init__(self, config, layer_idx: Optional[int] = None, dtype=torch.bfloat16):
        super().__init__()
        self.config, self.layer_idx = config, layer_idx
        self.hidden_size = config.hidden_size
        self.num_heads = config.num_attention_heads
        self.num_key_value_heads = config.num_key_value_heads
        self.head_dim = getattr(config, "head_dim", None) or config.hidden_size // config.num_attention_heads
        if getattr(config, "attention_bias", False):
            raise NotImplementedError("attention_bias=True is not supported by the B200 span kernels")
        if getattr(config, "pretraining_tp", 1) != 1:
            raise NotImplementedError("pretraining_tp > 1 (modules.py:44-59) is a single-device numerics emulation; use 1")
        q, kv, h = self.num_heads * self.head_dim, self.num_key_value_heads * self.head_dim, self.hidden_size
        self.q_proj = _Weight(q, h, dtype=dtype)
        self.k_proj = _Weight(kv, h, dtype=dtype)
        self.v_proj = _Weight(kv, h, dtype=dtype)
        self.o_proj = _Weight(h, q, dtype=dtype)


class LlamaMLPWeights(nn.Module):
    """Weights of HF ``LlamaMLP`` as built at modules.py:123."""

    def __init__(self, config, dtype=torch.bfloat16):
        super().__init__()
        if getattr(config, "mlp_bias", False):
            raise NotImplementedError("mlp_bias=True is not supported by the B200 span kernels")
        if getattr(config, "hidden_act", "silu") != "silu":
            raise NotImplementedError("only hidden_act='silu' is implemented")
        h, i = config.hidden_size, config.intermediate_size
        self.gate_proj = _Weight(i, h, dtype=dtype)
        self.up_proj = _Weight(i, h, dtype=dtype)
        self.down_proj = _Weight(h, i, dtype=dtype)


class OptimizedLlamaDecoderLayer(nn.Module):
    """Weights of reference ``OptimizedLlamaDecoderLayer`` (modules.py:117-184).

    ``fuse()`` packs q|k|v and gate|up into contiguous buffers on the parameters' device and re-points the
    parameters at views of them; it is re-run automatically when the parameters have been moved.
    """

    def __init__(self, config, layer_idx: int, dtype=torch.bfloat16):
        super().__init__()
        self.layer_idx = layer_idx
        self.hidden_size = config.hidden_size
        self.self_attn = OptimizedLlamaInferenceAttention(config, layer_idx, dtype=dtype)
        self.mlp = LlamaMLPWeights(config, dtype=dtype)
        self.input_layernorm = _Weight(config.hidden_size, dtype=dtype, ones=True)           # eps 1e-6: modules.py:124
        self.post_attention_layernorm = _Weight(config.hidden_size, dtype=dtype, ones=True)  # eps 1e-6: modules.py:125
        self._wqkv: Optional[torch.Tensor] = None
        self._wgu: Optional[torch.Tensor] = None
        self.w8 = None      # after quantize_w8(): {"qkv": (uint8 [n, k], fp32 [n]), "o": ..., "gu": ..., "down": ...}

    @torch.no_grad()
    def quantize_w8(self) -> None:
        """Replace the seven linear weights by row-wise absmax int8 + one fp32 scale per output row -- the weight format
        of the reference's optimised mode (bitsandbytes ``Linear8bitLt(has_fp16_weights=False)``, utils/model.py:102-108)
        -- computed on the GPU by ``dli_quantize_rows_w8``; the bf16 copies are released (the ``*.weight`` parameters
        become empty).  Decode then streams half the weight bytes (``dli_gemm_w8``)."""
        from distributed_llm_inference_b200 import _ffi

        if self.w8 is not None:
            return
        self.fuse()
        a, m = self.self_attn, self.mlp
        if not self._wqkv.is_cuda:
            raise RuntimeError("quantize_w8 runs on the GPU: move the block to CUDA first")
        out = {}
        with torch.cuda.device(self._wqkv.device):
            for name, w in (("qkv", self._wqkv), ("o", a.o_proj.weight.data), ("gu", self._wgu), ("down", m.down_proj.weight.data)):
                w = w.contiguous()
                n, k = w.shape
                q = torch.empty(n, k, dtype=torch.uint8, device=w.device)
                sc = torch.empty(n, dtype=torch.float32, device=w.device)
                _ffi.check(_ffi.lib.dli_quantize_rows_w8(w.data_ptr(), n, k, q.data_ptr(), sc.data_ptr(),
                                                         _ffi.current_stream(w.device)))
                out[name] = (q, sc)
            torch.cuda.synchronize(self._wqkv.device)
        self.w8 = out
        empty = torch.empty(0, dtype=self._wqkv.dtype, device=self._wqkv.device)
        for p in (a.q_proj.weight, a.k_proj.weight, a.v_proj.weight, a.o_proj.weight, m.gate_proj.weight, m.up_proj.weight,
                  m.down_proj.weight):
            p.data = empty
        self._wqkv = self._wgu = None

    def _is_fused(self) -> bool:
        a, m = self.self_attn, self.mlp
        if self._wqkv is None or self._wgu is None:
            return False
        q, kv = a.q_proj.weight.shape[0], a.k_proj.weight.shape[0]
        es = self._wqkv.element_size() * self._wqkv.shape[1]
        base = self._wqkv.data_ptr()
        ok = (a.q_proj.weight.data_ptr() == base and a.k_proj.weight.data_ptr() == base + q * es
              and a.v_proj.weight.data_ptr() == base + (q + kv) * es)
        gb = self._wgu.data_ptr()
        ok = ok and m.gate_proj.weight.data_ptr() == gb and \
            m.up_proj.weight.data_ptr() == gb + m.gate_proj.weight.shape[0] * es
        return ok

    @torch.no_grad()
    def fuse(self) -> None:
        if self.w8 is not None or self._is_fused():
            return
        a, m = self.self_attn, self.mlp
        wqkv = torch.cat([a.q_proj.weight.data, a.k_proj.weight.data, a.v_proj.weight.data], dim=0).contiguous()
        q, kv = a.q_proj.weight.shape[0], a.k_proj.weight.shape[0]
        a.q_proj.weight.data, a.k_proj.weight.data, a.v_proj.weight.data = wqkv[:q], wqkv[q:q + kv], wqkv[q + kv:]
        wgu = torch.cat([m.gate_proj.weight.data, m.up_proj.weight.data], dim=0).contiguous()
        i = m.gate_proj.weight.shape[0]
        m.gate_proj.weight.data, m.up_proj.weight.data = wgu[:i], wgu[i:]
        for p in (a.o_proj.weight, m.down_proj.weight, self.input_layernorm.weight, self.post_attention_layernorm.weight):
            if not p.data.is_contiguous():
                p.data = p.data.contiguous()
        self._wqkv, self._wgu = wqkv, wgu

    def forward(self, *a, **k):  # pragma: no cover
        raise RuntimeError("B200 span: layers are executed by LlamaBlock.forward (dli_span_forward), not one by one")
