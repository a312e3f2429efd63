"""CPU checks of oracle/quant_oracle.py (the restatement the int8 GPU tests compare against; unpinned): the two activation
planes are a partition of the row, the row-wise outlier rule stays within int8 noise of the column-wise rule of
bitsandbytes as restated, and both stay close to the unquantised product."""
import torch

from oracle import quant_oracle as qo


def _data(seed=3, m=12, n=96, k=384):
    g = torch.Generator().manual_seed(seed)
    w = 0.02 * torch.randn(n, k, generator=g)
    x = torch.randn(m, k, generator=g).clamp_(-4.5, 4.5)
    x[1, 17], x[5, 200], x[5, 17], x[9, 300] = 9.0, -7.5, 3.0, 11.0
    return x, w


def test_activation_planes_partition_the_row():
    x, _ = _data()
    c1, s1, c2, s2 = qo.quantize_act_rowwise(x, 5.0)
    out = x.abs() >= 5.0
    assert int(out.sum()) == 3
    assert torch.all(c1[out] == 0) and torch.all(c2[~out] == 0)
    assert torch.all(c2[out].abs() <= 127) and int(c2[9, 300]) == 127 and int(c2[5, 200]) == -127
    back = c1.float() * s1[:, None] + c2.float() * s2[:, None]
    assert (back - x).abs().max() <= 0.5 * float(s1.max()) + 1e-6        # outliers alone in their row: exact up to fp32
    n1, t1, n2, t2 = qo.quantize_act_rowwise(x, 0.0)                     # no threshold: one plane, the outlier sets the scale
    assert int((n2 != 0).sum()) == 0 and float(t2.max()) == 0.0 and abs(float(t1[9]) - 11.0 / 127.0) < 1e-7


def test_rowwise_outlier_rule_is_within_int8_noise_of_the_columnwise_rule():
    x, w = _data()
    q, s = qo.quantize_rowwise(w)
    exact = x @ qo.dequantize_rowwise(q, s).t()
    row, col, plain = (qo.linear_llm_int8_rowwise(x, q, s, 5.0), qo.linear_llm_int8(x, q, s, 5.0),
                       qo.linear_llm_int8_rowwise(x, q, s, 0.0))
    scale = exact.abs().max()
    assert (row - col).abs().max() <= 2e-3 * scale
    assert (row - exact).abs().max() <= 1.5e-2 * scale and (col - exact).abs().max() <= 1.5e-2 * scale
    # what the threshold buys: rows that hold an outlier lose resolution without it
    bad = [1, 5, 9]
    assert (plain[bad] - exact[bad]).abs().max() > (row[bad] - exact[bad]).abs().max()
