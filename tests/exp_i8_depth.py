"""Experiment (not a test): error of the two int8 modes at full depth.  7B-shaped span, 32 layers, B = 4: prefill 512
tokens (the quantised spans run chunks of more than 128 tokens on 16-bit activations) and 8 decode steps, compared at the
span output with the fp32 oracle on (a) the ORIGINAL weights -- what the mode costs in total -- and (b) the row-wise
dequantised weights -- what the activations' share is.  Printed: max | mean error / max-abs of the oracle's output."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
from distributed_llm_inference_b200.utils import synthetic as syn
from oracle import quant_oracle as qo
from oracle import span_oracle as so
from tests.test_deep_gpu import _gpu_span, _oracle_on

DEV = "cuda"
torch.backends.cuda.matmul.allow_tf32 = False
name, n_layers, B, T, n_dec = os.environ.get("EXP_SHAPE", "7b"), int(os.environ.get("EXP_LAYERS", "32")), 4, 512, 8
shp = syn.SHAPES[name]
ids = list(range(n_layers))
W = [syn.make_layer_weights(shp, g, device=DEV) for g in ids]


def dequantised(w):
    d = dict(w)
    for k in syn.LINEARS:
        d[k] = qo.dequantize_rowwise(*qo.quantize_rowwise(w[k])).to(torch.bfloat16)
    return d


# q|k|v and gate|up are quantised as fused matrices, but row-wise: per-row scales do not depend on the fusion
Wd = [dequantised(w) for w in W]
oracles = {"original weights": _oracle_on(DEV, shp, W), "dequantised weights": _oracle_on(DEV, shp, Wd)}
del Wd
spans = {"bf16": _gpu_span(shp, W, ids),
         "int8 weights, bf16 activations": _gpu_span(shp, W, ids).quantize_w8(activations="bf16"),
         "LLM.int8 (threshold 5)": _gpu_span(shp, W, ids).quantize_w8(activations="int8", threshold=5.0),
         "int8 x int8, no threshold": _gpu_span(shp, W, ids).quantize_w8(activations="int8", threshold=0.0)}
del W
caches = {k: PartialLlamaSinkCache(1 << 20, 4, capacity_tokens=B * (T + 64) + 1024) for k in spans}
ocache = {k: so.SinkCacheOracle(1 << 20, 4) for k in oracles}
pos, rows = 0, {k: [] for k in spans}
for step, L in enumerate([T] + [1] * n_dec):
    x = syn.make_hidden(shp, B, L, seed=80 + step, device=DEV)
    cp = torch.arange(pos, pos + L, device=DEV)
    refs = {k: o.forward("g", x.float(), past_key_value=ocache[k], cache_position=cp)[0] for k, o in oracles.items()}
    for k, blk in spans.items():
        got = blk("g", x, past_key_value=caches[k], cache_position=cp)[0].float()
        rows[k].append({r: ((got - ref).abs().max().item() / ref.abs().max().item(),
                            (got - ref).abs().mean().item() / ref.abs().max().item()) for r, ref in refs.items()})
    pos += L
print(f"{name} x {n_layers} layers, B={B}: span output, error / max-abs of the fp32 oracle (max | mean)")
for k in spans:
    for r in oracles:
        pre = rows[k][0][r]
        dec = [rows[k][i][r] for i in range(1, n_dec + 1)]
        print(f"  {k:32s} vs oracle on {r:19s}: prefill {pre[0]:.2e}|{pre[1]:.2e}   decode steps max {max(d[0] for d in dec):.2e} "
              f"mean {sum(d[1] for d in dec) / len(dec):.2e}")
