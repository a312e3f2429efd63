"""Decode attention without caller scratch (dli_decode_attention_static).

Who needs it: code that is given input and output buffers and may not allocate — a plug-in living inside a host
framework that already owns all of HBM, or a serving loop where a cudaMalloc (device-wide synchronisation, or an
out-of-memory failure when the pool is full) in the middle of a request is not acceptable.  The entry point keeps
its split-K partials in an 8 MiB array compiled into the library instead.

The test drives it through the C ABI twice in a row, with different inputs, and compares each result with the
ordinary entry point (caller-provided scratch) on the same inputs AND with the attention arithmetic of the
reference (modules.py:87-105) evaluated in fp32 on the host, so an error shared by both entry points cannot hide;
the second call would expose anything that wrongly survives in the shared array.  It also checks that a problem too big for the array is refused with a
message rather than run.  The file name makes pytest collect it last, so a failure here cannot hide the other tests.
"""
import ctypes as C

import pytest
import torch

from distributed_llm_inference_b200 import _ffi
from tests import gpu_ops as ops
from tests.test_ops_gpu import _attn_ref

pytestmark = pytest.mark.gpu


def _static(q, kv, layer, B, H, KV, hd, max_ctx):
    out = torch.empty(B, H * hd, dtype=torch.bfloat16, device=q.device)
    d = kv.desc(max_ctx)
    rc = _ffi.lib.dli_decode_attention_static(q.data_ptr(), out.data_ptr(), C.byref(d), layer, B, H, KV, hd, ops.stream())
    return rc, out


@pytest.mark.parametrize("H,KV,hd", [(8, 2, 64), (4, 4, 128)])
def test_static_scratch_matches_caller_scratch_twice(H, KV, hd):
    g = torch.Generator().manual_seed(1234)
    for call, (B, T) in enumerate([(3, 700), (2, 90)]):          # second call: smaller problem, other data
        kv = ops.PagedKV(1, B, KV, hd, T + 3, seed=call)
        K = torch.randn(B, KV, T, hd, generator=g).to(torch.bfloat16).cuda()
        V = torch.randn(B, KV, T, hd, generator=g).to(torch.bfloat16).cuda()
        kv.fill(0, K, V)
        kv.seq_lens.fill_(T - 1)
        q = torch.randn(B, H * hd, generator=g).to(torch.bfloat16).cuda()
        want = ops.decode_attention(q, kv, 0, B, H, KV, hd, max_ctx=T)
        rc, got = _static(q, kv, 0, B, H, KV, hd, T)
        assert rc == 0, _ffi.lib.dli_last_error()
        torch.cuda.synchronize()
        msg, err, scale = ops.err_report(f"static decode attention call {call}", got, want)
        assert err <= 8e-3 * scale + 1e-3, msg                    # same bound as test_decode_attention
        ref = _attn_ref(q.view(B, H, 1, hd).float().cpu(), K.float().cpu(), V.float().cpu(), False, T - 1)[:, 0]
        msg, err, scale = ops.err_report(f"static decode attention call {call} vs fp32 reference math", got, ref)
        assert err <= 8e-3 * scale + 1e-3, msg


def test_static_scratch_refuses_oversized_problem():
    B, H, KV, hd, T = 64, 64, 1, 128, 4096                        # 64*64*5 splits*130*4 B = 10.6 MB > 8 MiB
    kv = ops.PagedKV(1, B, KV, hd, 16)                            # never read: the call must fail before launching
    q = torch.zeros(B, H * hd, dtype=torch.bfloat16, device="cuda")
    rc, _ = _static(q, kv, 0, B, H, KV, hd, T)
    assert rc == -4
    msg = _ffi.lib.dli_last_error()
    assert b"built-in array" in msg and b"dli_decode_attention" in msg
