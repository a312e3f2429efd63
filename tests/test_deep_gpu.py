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


def _oracle_bf16_on(device, shp, weights, rope="float"):
    """The reference's arithmetic executed in bf16 (what its own bf16 run computes: every op rounds to bf16,
    modules.py:61-63,90,96-97,112) -- the yardstick for how much error the activation dtype itself costs."""
    shape = so.SpanShape(shp.hidden_size, shp.intermediate_size, shp.num_attention_heads,
                         shp.num_key_value_heads, shp.head_dim, shp.rope_theta)
    layers = [so.LayerWeights(*(w[k].to(device).to(torch.bfloat16) for k in syn.LINEARS + syn.NORMS)) for w in weights]
    return so.OracleSpan(shape, layers, "eager", rope)


@pytest.mark.parametrize("shape_name,n_layers,B,T,n_dec", [("7b", 32, 4, 512, 8), ("70b", 10, 8, 512, 8)])
def test_full_depth_span_matches_fp32_oracle_on_gpu(shape_name, n_layers, B, T, n_dec):
    """Whole 7B span (32 layers) and one 70B pipeline stage (10 layers): prefill 512 + 8 decode steps, every layer
    boundary compared with the fp32 oracle.  Error grows with depth and width (every layer rounds its GEMM operands to
    bf16), so a fixed percentage is not a meaningful bound at depth 32; the yardstick is the reference's OWN arithmetic
    executed in bf16 (the dtype the north star names) against the same fp32 oracle: at every depth this library's
    max-abs and mean-abs errors must not exceed it, at the span output the mean-abs error must be <= 0.7 x it
    (measured: 0.45-0.52 x, fp32 residual stream and fp32 intermediates inside the fused regions), and the max-abs
    error stays under 6 % of the tensor's max-abs.  The table of error against depth is printed (DESIGN.md quotes it)."""
    torch.backends.cuda.matmul.allow_tf32 = False
    shp = syn.SHAPES[shape_name]
    ids = list(range(n_layers))
    W = [syn.make_layer_weights(shp, g, device=DEV) for g in ids]
    blk = _gpu_span(shp, W, ids)
    osp = _oracle_on(DEV, shp, W)
    obf = _oracle_bf16_on(DEV, shp, W)
    del W
    gc = PartialLlamaSinkCache(1 << 20, 4, capacity_tokens=B * (T + 64) + 1024)
    oc, bc = so.SinkCacheOracle(1 << 20, 4), so.SinkCacheOracle(1 << 20, 4)
    rows, failures = [], []
    pos = 0

    def rel_err(a, r):
        scale = r.abs().max().item()
        return (a - r).abs().max().item() / scale, (a - r).abs().mean().item() / scale

    for step, L in enumerate([T] + [1] * n_dec):
        x = syn.make_hidden(shp, B, L, seed=80 + step, device=DEV)
        cp = torch.arange(pos, pos + L, device=DEV)
        ref = osp.forward("g", x.float(), past_key_value=oc, cache_position=cp, output_hidden_states=True)[1]
        bfr = obf.forward("g", x, past_key_value=bc, cache_position=cp, output_hidden_states=True)[1]
        got = blk("g", x, past_key_value=gc, cache_position=cp, output_hidden_states=True)[1]
        assert len(ref) == len(got) == len(bfr) == n_layers + 1
        assert all(torch.isfinite(g.float()).all() for g in got)
        ours = [rel_err(got[i].float(), ref[i]) for i in range(n_layers + 1)]
        theirs = [rel_err(bfr[i].float(), ref[i]) for i in range(n_layers + 1)]
        rows.append((step, L, ours, theirs))
        for d in range(1, n_layers + 1):
            if ours[d][0] > theirs[d][0] or ours[d][1] > theirs[d][1] or ours[d][0] > 6e-2:
                failures.append(f"step {step} depth {d}: max {ours[d][0]:.3e} mean {ours[d][1]:.3e} vs the reference in bf16 "
                                f"max {theirs[d][0]:.3e} mean {theirs[d][1]:.3e}")
        if ours[-1][1] > 0.7 * theirs[-1][1]:
            failures.append(f"step {step} output: ours max {ours[-1][0]:.3e} mean {ours[-1][1]:.3e} vs the reference in bf16 "
                            f"max {theirs[-1][0]:.3e} mean {theirs[-1][1]:.3e}")
        pos += L
    print(f"\n{shape_name} x {n_layers} layers, B={B}: error / max-abs of the fp32 oracle, by depth (max | mean)")
    depths = sorted({1, 2, 4, 8, 10, 16, 24, n_layers} & set(range(n_layers + 1)))
    print("  depth                      " + " ".join(f"{d:>17d}" for d in depths))
    for step, L, ours, theirs in (rows[0], rows[1], rows[-1]):
        print(f"  q_len {L:<4d} this library     " + " ".join(f"{ours[d][0]:8.2e}|{ours[d][1]:8.2e}" for d in depths))
        print(f"  q_len {L:<4d} reference in bf16" + " ".join(f"{theirs[d][0]:8.2e}|{theirs[d][1]:8.2e}" for d in depths))
    assert not failures, "; ".join(failures[:6])


def test_run_to_run_difference_is_bounded():
    """Two generations fed identical inputs through the same span: the only source of difference is the order of the
    ``red.global.add`` reductions (split-K partials, residual adds).  Bound: four bf16 ulps (2^-6) of the tensor's
    max-abs after prefill and after each of 4 decode steps (measured on B200: 0 after the prefill -- the prefill GEMMs
    do not split K -- then 3.2e-3 .. 6.7e-3, with 46-60 % of the elements bit-identical)."""
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
    assert max(worst) <= 2.0 ** -6
