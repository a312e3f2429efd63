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
half exactly (``quantize_rowwise`` below is what ``dli_quantize_rows_w8`` must reproduce bit for bit); activations are
either kept in bf16 (``linear_w8`` is that arithmetic in fp32) or quantised per token row (``quantize_act_rowwise`` /
``linear_llm_int8_rowwise`` are what ``dli_quantize_act_i8`` / ``dli_gemm_i8`` must reproduce); ``linear_llm_int8`` is
the reference's column-wise outlier rule for comparison.
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


def quantize_act_rowwise(x: torch.Tensor, threshold: float = 5.0):
    """LLM.int8 activation quantisation of ``x [m, k]`` as ``dli_quantize_act_i8`` does it, per token row:
    -> (c1 int8 [m, k], s1 fp32 [m], c2 int8 [m, k], s2 fp32 [m]).  Entries with ``|x| >= threshold`` are outliers: they
    are 0 in c1 and do not enter s1 = absmax of the remaining entries / 127 (bitsandbytes' row statistics likewise leave
    them out [recall]); c1 = round(x * (127 / absmax)) clamped to +-127, ties to even.  The outliers form the second
    plane, quantised the same way with s2 = absmax over the outliers / 127 (c2 is 0 elsewhere)."""
    x = x.float()
    out = (x.abs() >= threshold) if threshold > 0 else torch.zeros_like(x, dtype=torch.bool)

    def plane(mask):
        absmax = torch.where(mask, x.abs(), torch.zeros_like(x)).amax(dim=1)
        inv = torch.where(absmax > 0, torch.full_like(absmax, 127.0) / absmax, torch.zeros_like(absmax))
        codes = torch.round(x * inv[:, None]).clamp_(-127, 127)
        return torch.where(mask, codes, torch.zeros_like(codes)).to(torch.int8), absmax / 127.0

    c1, s1 = plane(~out)
    c2, s2 = plane(out)
    return c1, s1, c2, s2


def linear_llm_int8_rowwise(x: torch.Tensor, q: torch.Tensor, scale: torch.Tensor, threshold: float = 5.0) -> torch.Tensor:
    """What ``dli_gemm_i8`` computes: both activation planes int8 x int8 in exact integers, each times the token's plane
    scale, summed, times the feature's weight scale."""
    c1, s1, c2, s2 = quantize_act_rowwise(x, threshold)
    qd = q.double()
    res = (c1.double() @ qd.t()).float() * s1[:, None] + (c2.double() @ qd.t()).float() * s2[:, None]
    return res * scale[None, :].float()


def linear_llm_int8(x: torch.Tensor, q: torch.Tensor, scale: torch.Tensor, threshold: float = 5.0) -> torch.Tensor:
    """bitsandbytes MatMul8bitLt.forward for has_fp16_weights=False, restated [recall]: a value at or above the threshold
    makes its COLUMN an outlier column for every row of the batch; those columns are multiplied in 16 bits against the
    dequantised weight columns, the rest int8 x int8 -> int32 with row-wise activation scales (row absmax over the
    entries below the threshold).  ``linear_llm_int8_rowwise`` differs in the outlier columns only: 8 bits relative to
    the row's largest outlier there (and to the regular scale for the other rows' entries), 16 bits here."""
    x = x.float()
    c1, s1, _, _ = quantize_act_rowwise(x, threshold)
    cols = (x.abs() >= threshold).any(dim=0) if threshold > 0 else torch.zeros(x.shape[1], dtype=torch.bool)
    c1 = c1.clone()
    c1[:, cols] = 0
    res = (c1.double() @ q.double().t()).float() * s1[:, None]
    if cols.any():
        res = res + x[:, cols] @ q.float()[:, cols].t()
    return res * scale[None, :].float()
