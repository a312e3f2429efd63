"""Row-wise int8 weight quantisation — TEST INFRASTRUCTURE, NOT PRODUCT CODE (same rules as span_oracle.py).

The reference's optimised mode (utils/model.py:93-123) replaces every ``nn.Linear`` by bitsandbytes
``Linear8bitLt(threshold=5.0, has_fp16_weights=False)``.  bitsandbytes 0.42.0 (poetry.lock:49-50) is not installed and
cannot be (no network, no sm_100 build), so its algorithm is restated here from its published source [recall]:

  weights      ``Int8Params.cuda()`` -> ``double_quant``: CB[r, :] = round(W[r, :] * 127 / absmax_r) as int8,
               SCB[r] = absmax_r (row-wise absmax over the input dimension).
  activations  LLM.int8: per token row, values with |x| >= threshold mark their COLUMN as an outlier column; those
               columns are multiplied in fp16 against the dequantised weight columns, the rest is quantised row-wise to
               int8 (SCA = absmax of the remaining entries), multiplied in int32 and dequantised by SCA * SCB / 127^2.

Parity pinning: NONE — there is no bitsandbytes here to run, and the reference has no fixtures for this mode, so this
file is *unpinned* ("parity unpinned" applies to everything that depends on it).  The B200 path implements the WEIGHT
half exactly (``quantize_rowwise`` below is what ``dli_quantize_rows_w8`` must reproduce bit for bit) and keeps
activations in bf16; ``linear_w8`` is its arithmetic in fp32, ``linear_llm_int8`` the reference's for comparison.
"""
from __future__ import annotations

from typing import Tuple

import torch


def quantize_rowwise(w: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (q int8 [n, k], scale fp32 [n]) with scale = absmax / 127 and q = round(w / scale) (ties to even, as the GPU's
    cvt.rni), zero rows mapping to scale 0."""
    w = w.float()
    absmax = w.abs().amax(dim=1)
    # tensor / tensor: a correctly rounded fp32 division (scalar / tensor is reciprocal-then-multiply in torch)
    inv = torch.where(absmax > 0, torch.full_like(absmax, 127.0) / absmax, torch.zeros_like(absmax))
    q = torch.round(w * inv[:, None]).clamp_(-127, 127).to(torch.int8)
    return q, absmax / 127.0


def dequantize_rowwise(q: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    return q.float() * scale[:, None].float()


def linear_w8(x: torch.Tensor, q: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    """What the B200 int8-weight GEMM computes: exact integer weights times the activations, scaled per output row."""
    return (x.float() @ q.float().t()) * scale[None, :].float()


def linear_llm_int8(x: torch.Tensor, q: torch.Tensor, scale: torch.Tensor, threshold: float = 5.0) -> torch.Tensor:
    """bitsandbytes MatMul8bitLt.forward for has_fp16_weights=False, restated [recall]: outlier columns in 16 bits, the
    rest int8 x int8 -> int32 with row-wise activation scales."""
    x = x.float()
    outlier_cols = (x.abs() >= threshold).any(dim=0) if threshold > 0 else torch.zeros(x.shape[1], dtype=torch.bool)
    xin = x.clone()
    xin[:, outlier_cols] = 0
    sca = xin.abs().amax(dim=1).clamp_min(1e-30)
    ca = torch.round(xin * (127.0 / sca)[:, None]).clamp_(-127, 127)
    out = (ca @ q.float().t()) * (sca[:, None] * (scale * 127.0)[None, :] / (127.0 * 127.0))
    if outlier_cols.any():
        wd = dequantize_rowwise(q, scale)[:, outlier_cols]
        out = out + x[:, outlier_cols] @ wd.t()
    return out
