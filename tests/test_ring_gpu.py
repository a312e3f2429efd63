"""Pipeline stage steps (dli_stage_decode_step) on ONE GPU: the stage is first and last at once, so the ring closes on
itself -- the same entry kernel (flag wait / embedding gather), exit kernel (store, fence, counter, flag release), head
(final RMSNorm -> lm_head -> argmax) and round graph that run between GPUs, checked against the plain span path and the
sequential client.  The multi-GPU run of the same protocol is checked by ``bench.py``'s ``pipeline_parity`` record
(N-stage output against the same layers as one span and against the oracle) on every N > 1 launch."""

import pytest
import torch

from distributed_llm_inference_b200 import _ffi
from distributed_llm_inference_b200.client import SequentialClient
from distributed_llm_inference_b200.models.llama import PartialLlamaSinkCache
from distributed_llm_inference_b200.parallel.pipeline import PipelineStage
from distributed_llm_inference_b200.utils import synthetic as syn
from tests.test_span_gpu import gpu_make_span

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_lm_head_argmax_matches_torch():
    """Client tail as one C call against torch arithmetic on the same bf16 values (RMSNorm in fp32, rounded to bf16 as
    the GEMM operand, fp32 accumulation); tokens are the argmax of the logits the kernel itself produced."""
    g = torch.Generator().manual_seed(5)
    for B, h, V in [(1, 512, 32000), (16, 4096, 32000), (5, 768, 1024)]:
        x = torch.randn(B, h, generator=g).to(torch.bfloat16).to(DEV)
        nw = (1 + 0.1 * torch.randn(h, generator=g)).to(torch.bfloat16).to(DEV)
        W = (0.02 * torch.randn(V, h, generator=g)).to(torch.bfloat16).to(DEV)
        xn = torch.empty(B, h, dtype=torch.bfloat16, device=DEV)
        logits = torch.empty(B, V, dtype=torch.float32, device=DEV)
        toks = torch.empty(B, dtype=torch.int32, device=DEV)
        _ffi.check(_ffi.lib.dli_lm_head_argmax(x.data_ptr(), 0, nw.data_ptr(), 1e-6, W.data_ptr(), V, h, B, xn.data_ptr(),
                                               logits.data_ptr(), toks.data_ptr(), _ffi.current_stream()))
        x32 = x.float()
        ref_n = (nw.float() * (x32 * torch.rsqrt(x32.pow(2).mean(-1, keepdim=True) + 1e-6))).to(torch.bfloat16)
        ref = ref_n.float() @ W.float().t()
        scale = ref.abs().max().item()
        assert (logits - ref).abs().max().item() <= 2e-3 * scale + 1e-4      # one bf16 ulp of xn can flip in the norm
        assert torch.equal(toks.long(), logits.argmax(-1))


def _client_and_spans(shp, layer_splits):
    cw = {k: v.to(DEV) for k, v in syn.make_decidable_client_weights(shp).items()}
    spans = [gpu_make_span(shp, syn.make_span_weights(shp, list(range(a, b))), "eager", "float", first_layer=a)
             for a, b in layer_splits]
    return cw, spans


def test_token_ring_on_one_gpu_matches_sequential_client():
    """Two micro-batches (different prompts) decode 12 tokens through the token ring -- embedding gather in the entry
    kernel, head + argmax closing the step, token ids fed back through the device-resident inbox, whole rounds replayed
    as one chained graph -- and produce exactly the tokens the sequential client produces step by step."""
    shp = syn.SHAPES["test_hd128"]
    cw, (span,) = _client_and_spans(shp, [(0, 4)])
    B, T, n_new = 3, 40, 12
    prompts = [syn.make_prompt(shp, B, T, seed=200 + m, device=DEV) for m in range(2)]
    # reference run: sequential client on its own cache
    ref_cache = PartialLlamaSinkCache(4096, 4, capacity_tokens=4096)
    ref_client = SequentialClient(cw["embed_tokens"], cw["norm"], cw["lm_head"], [span], [ref_cache])
    want = [ref_client.greedy(f"ref{m}", prompts[m], n_new + 1)[0] for m in range(2)]       # [B, n_new + 1]
    # ring run: prefill through the same client, first token from its head, then the ring feeds itself
    cache = PartialLlamaSinkCache(4096, 4, capacity_tokens=4096)
    client = SequentialClient(cw["embed_tokens"], cw["norm"], cw["lm_head"], [span], [cache])
    gids = ["mb0", "mb1"]
    first_tok = []
    for m, gid in enumerate(gids):
        client.step(gid, prompts[m], 0)
        first_tok.append(client.last_tokens.clone())
        assert torch.equal(first_tok[m].long(), want[m][:, 0])
    stage = PipelineStage(span, cache, 0, 1, torch.device(DEV),
                          client={"embed_tokens": cw["embed_tokens"], "norm": cw["norm"], "lm_head": cw["lm_head"]})
    assert stage.enable_graphs(gids, T + n_new + 4, batch=B, ring="tokens") == "local"
    got = [[], []]
    xs = first_tok
    for step in range(n_new):
        if step == 3:
            assert stage.capture_round(gids, ring="tokens")
        out = stage.run_round(gids, xs, B, 1, T + step, ring="tokens")
        stage.drain()
        for m in range(2):
            got[m].append(out[m].clone())
        xs = None                                        # from now on the ring feeds itself
    stage.check_errors()
    for m in range(2):
        toks = torch.stack(got[m], dim=1).long()
        assert torch.equal(toks, want[m][:, 1:]), (m, toks.tolist(), want[m][:, 1:].tolist())
    assert cache.get_seq_length(0, "mb0") == T + n_new


def test_hidden_ring_on_one_gpu_matches_eager_steps():
    """Hidden-state ring closed on itself (exit cast stores into the stage's own inbox, flag wait with lag 1) against
    the same steps run eagerly through LlamaBlock.forward on a second copy of the generation."""
    shp = syn.SHAPES["test_gqa"]
    _, (span,) = _client_and_spans(shp, [(0, 3)])
    B, T, n_new = 2, 33, 6
    cache = PartialLlamaSinkCache(4096, 4, capacity_tokens=4096)
    x0 = syn.make_hidden(shp, B, T, seed=1, device=DEV)
    for gid in ("ring", "eager"):
        span(gid, x0, past_key_value=cache)
    stage = PipelineStage(span, cache, 0, 1, torch.device(DEV))
    stage.enable_graphs(["ring"], T + n_new + 2, batch=B, ring="hidden")
    x = syn.make_hidden(shp, B, 1, seed=2, device=DEV)
    xe, xs = x.clone(), [x.clone()]
    for step in range(n_new):
        out = stage.run_round(["ring"], xs, B, 1, T + step)
        stage.drain()
        xe = span("eager", xe, past_key_value=cache, cache_position=torch.tensor([T + step], device=DEV))[0]
        scale = xe.float().abs().max().item()
        assert (out[0].float() - xe.float()).abs().max().item() <= 2e-2 * scale, step
        xs = None
    stage.check_errors()


def test_flag_wait_times_out_instead_of_hanging():
    """A wait whose producer never signals gives up after the time-out and records it in the step state."""
    flag = torch.zeros(16, dtype=torch.int32, device=DEV)
    state = torch.zeros(4, dtype=torch.int32, device=DEV)
    _ffi.check(_ffi.lib.dli_flag_wait(flag.data_ptr(), 5, state.data_ptr(), 50, _ffi.current_stream()))
    torch.cuda.synchronize()
    assert state.tolist()[2] == 1
    _ffi.check(_ffi.lib.dli_flag_wait(flag.data_ptr(), 5, state.data_ptr(), 60000, _ffi.current_stream()))   # sticky: returns at once
    torch.cuda.synchronize()
    flag[0] = 7
    state.zero_()
    _ffi.check(_ffi.lib.dli_flag_wait(flag.data_ptr(), 5, state.data_ptr(), 50, _ffi.current_stream()))
    torch.cuda.synchronize()
    assert state.tolist()[2] == 0
