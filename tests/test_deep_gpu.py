"""Full-depth parity at the bench shapes: the fp32 oracle itself run ON THE GPU.

The CPU oracle is affordable only for a few full-width layers (tests/test_fullsize_gpu.py), which leaves error growth
over the 32 layers of the 7B span and the 10 layers of a 70B stage unmeasured, and lets an error shared by two sides of
a self-consistency property pass.  ``oracle/span_oracle.py`` is device-agnostic torch: here it runs in fp32 (TF32 off)
on the same GPU, on the same bf16-representable weights, for the whole span -- a 512-token prefill and 8 cached decode
steps -- and every layer's input and the span output are compared (``output_hidden_states``).  The GPU run of the oracle
is first cross-validated against its CPU run on a small shape, so the checker is the pinned oracle, not a new one.

Also here: the bound on run-to-run differences.  Split-K partial sums and residual adds are reduced in L2 with
``red.global.add.f32`` whose order varies between runs, so outputs are reproducible to fp32 rounding of the sums, not
bit for bit; the test runs the same prefill + decode steps on two generations and bounds what that costs.
"""
import numpy as np
import pytest
import torch

from distributed_llm_inference_b200.models.llama import LlamaBlock, PartialLlamaSinkCache
from distributed_llm_inference_b200.utils import synthetic as syn
from oracle import span_oracle as so

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _oracle_on(device, shp, weights, rope="float"):
    shape = so.SpanShape(shp.hidden_size, shp.intermediate_size, shp.num_attention_heads,
                         shp.num_key_value_heads, shp.head_dim, shp.rope_theta)
    layers = [so.LayerWeights(*(w[k].to(device).float() for k in syn.LINEARS + syn.NORMS)) for w in weights]
    return so.OracleSpan(shape, layers, "eager", rope)


def _gpu_span(shp, weights, layer_ids, rope="float"):
    with torch.device(DEV):
        blk = LlamaBlock(shp.to_hf_config("eager"), list(layer_ids), rope=rope)
    for j, w in enumerate(weights):
        blk.load_state_dict({f"layers.{j}.{n}.weight": t for n, t in w.items()}, strict=False)
    return blk


def test_gpu_oracle_equals_cpu_oracle():
    """The checker used below (oracle in fp32 on the GPU) against the pinned CPU oracle, small shape, with cache."""
    torch.backends.cuda.matmul.allow_tf32 = False
    shp = syn.SHAPES["test_gqa"]
    W = syn.make_span_weights(shp, [0, 1, 2], norm_jitter=0.1)
    cpu, gpu = _oracle_on("cpu", shp, W), _oracle_on(DEV, shp, W)
    cc, gc = so.SinkCacheOracle(4096, 4), so.SinkCacheOracle(4096, 4)
    pos = 0
    for step, L in enumerate([33, 1, 1]):
        x = syn.make_hidden(shp, 2, L, seed=60 + step).float()
        cp = torch.arange(pos, pos + L)
        a = cpu.forward("g", x, past_key_value=cc, cache_position=cp, output_hidden_states=True)
        b = gpu.forward("g", x.to(DEV), past_key_value=gc, cache_position=cp.to(DEV), output_hidden_states=True)
        for i, (u, v) in enumerate(zip(a[1], b[1])):
            d = (u - v.cpu()).abs().max().item()
            assert d <= 2e-5 * max(1.0, u.abs().max().item()), (step, i, d)
        pos += L


@pytest.mark.parametrize("shape_name,n_layers,B,T,n_dec", [("7b", 32, 4, 512, 8), ("70b", 10, 8, 512, 8)])
def test_full_depth_span_matches_fp32_oracle_on_gpu(shape_name, n_layers, B, T, n_dec):
    """Whole 7B span (32 layers) and one 70B pipeline stage (10 layers): prefill 512 + 8 decode steps, every layer
    boundary compared.  Bound: max-abs error <= 3 % of the tensor's max-abs at every depth (the span-level tolerance of
    tests/test_span_gpu.py); the table of error against depth is printed (DESIGN.md quotes it)."""
    torch.backends.cuda.matmul.allow_tf32 = False
    shp = syn.SHAPES[shape_name]
    ids = list(range(n_layers))
    W = [syn.make_layer_weights(shp, g, device=DEV) for g in ids]
    blk = _gpu_span(shp, W, ids)
    osp = _oracle_on(DEV, shp, W)
    del W
    gc = PartialLlamaSinkCache(1 << 20, 4, capacity_tokens=B * (T + 64) + 1024)
    oc = so.SinkCacheOracle(1 << 20, 4)
    rows = []
    pos = 0
    for step, L in enumerate([T] + [1] * n_dec):
        x = syn.make_hidden(shp, B, L, seed=80 + step, device=DEV)
        cp = torch.arange(pos, pos + L, device=DEV)
        ref = osp.forward("g", x.float(), past_key_value=oc, cache_position=cp, output_hidden_states=True)[1]
        got = blk("g", x, past_key_value=gc, cache_position=cp, output_hidden_states=True)[1]
        assert len(ref) == len(got) == n_layers + 1
        rel = []
        for i in range(n_layers + 1):
            r, g = ref[i], got[i].float()
            assert torch.isfinite(g).all()
            scale = r.abs().max().item()
            rel.append(((g - r).abs().max().item() / scale, (g - r).abs().mean().item() / scale))
        rows.append((step, L, rel))
        worst = max(m for m, _ in rel)
        assert worst <= 3e-2, f"{shape_name} step {step}: max-abs error {worst:.3e} of scale at depth {int(np.argmax([m for m, _ in rel]))}"
        assert rel[-1][1] <= 6e-3, f"{shape_name} step {step}: mean-abs error {rel[-1][1]:.3e} of scale at the span output"
        pos += L
    print(f"\n{shape_name} x {n_layers} layers, B={B}: max-abs error / max-abs of the fp32 oracle, by depth")
    depths = sorted({0, 1, 2, 4, 8, 10, 16, 24, n_layers} & set(range(n_layers + 1)))
    print("  depth      " + " ".join(f"{d:>9d}" for d in depths))
    for step, L, rel in (rows[0], rows[1], rows[-1]):
        print(f"  q_len {L:<5d}" + " ".join(f"{rel[d][0]:9.2e}" for d in depths))


def test_run_to_run_difference_is_bounded():
    """Two generations fed identical inputs through the same span: the only source of difference is the order of the
    ``red.global.add`` reductions (split-K partials, residual adds).  Bound: two bf16 ulps (2^-7) of the tensor's
    max-abs after prefill and after each of 4 decode steps; typically the vast majority of elements are identical."""
    shp = syn.SHAPES["7b"]
    ids = [0, 1, 2, 3]
    W = [syn.make_layer_weights(shp, g, device=DEV) for g in ids]
    blk = _gpu_span(shp, W, ids)
    cache = PartialLlamaSinkCache(1 << 20, 4, capacity_tokens=2 * 16 * (256 + 64))
    B, T = 16, 256
    x = syn.make_hidden(shp, B, T, seed=3, device=DEV)
    outs = {g: blk(g, x, past_key_value=cache)[0].float() for g in ("a", "b")}
    worst, same = [], []
    for step in range(5):
        a, b = outs["a"], outs["b"]
        scale = a.abs().max().item()
        worst.append((a - b).abs().max().item() / scale)
        same.append((a == b).float().mean().item())
        xt = syn.make_hidden(shp, B, 1, seed=200 + step, device=DEV)
        cp = torch.tensor([T + step], device=DEV)
        outs = {g: blk(g, xt, past_key_value=cache, cache_position=cp)[0].float() for g in ("a", "b")}
    print(f"\nrun-to-run: max difference / scale per step {['%.2e' % w for w in worst]}, identical elements "
          f"{['%.4f' % s for s in same]}")
    assert max(worst) <= 2.0 ** -7
