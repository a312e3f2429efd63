"""ctypes binding of libdli_b200.so (the C ABI declared in include/dli_b200.h).

This is the stub a maintainer of the reference would add to call the sm_100a kernels from its Python
modules (the reference has no FFI of its own).  Nothing here touches torch types:
pointers travel as integers (``tensor.data_ptr()``), streams as ``torch.cuda.current_stream().cuda_stream``.

There is deliberately NO fallback: if the shared library is missing, importing this module fails with
instructions to build it; if a call fails, ``DliError`` carries ``dli_last_error()``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# DLI_LIB selects another build of the same sources (the DLI_TRACE=1 timeline build used by tests/exp_trace.py)
LIB_PATH = os.environ.get("DLI_LIB") or os.path.join(_HERE, "csrc", "libdli_b200.so")

DLI_OK = 0
EPI_STORE_F32, EPI_ADD_F32, EPI_STORE_BF16, EPI_SILU_MUL_BF16 = 0, 1, 2, 3

# every symbol include/dli_b200.h declares (tests check that the library exports all of them)
EXPORTS = (
    "dli_version", "dli_last_error", "dli_rmsnorm", "dli_gemm", "dli_rope_kv_append",
    "dli_decode_attention_scratch_bytes", "dli_decode_attention", "dli_prefill_attention", "dli_silu_mul",
    "dli_cast_f32_bf16", "dli_span_create", "dli_span_destroy", "dli_span_bind_weights",
    "dli_span_workspace_bytes", "dli_span_forward", "dli_last_launch_count", "dli_span_profile",
    "dli_span_profile_read", "dli_decode_attention_static", "dli_kv_evict", "dli_span_decode_step",
    "dli_stage_decode_step", "dli_flag_wait", "dli_lm_head_argmax",
    "dli_span_forward_masked", "dli_prefill_attention_masked", "dli_decode_attention_masked",
    "dli_quantize_rows_w8", "dli_dequantize_rows_w8", "dli_gemm_w8", "dli_span_bind_weights_w8",
    "dli_span_dequant_scratch_bytes", "dli_span_set_dequant_scratch",
    "dli_act_i8_bytes", "dli_act_i8_carve", "dli_quantize_act_i8", "dli_gemm_i8",
    "dli_span_act_i8_scratch_bytes", "dli_span_set_act_i8",
)


class DliError(RuntimeError):
    pass


class KvDesc(C.Structure):
    _fields_ = [("k_pages", C.c_void_p), ("v_pages", C.c_void_p), ("layer_stride_bytes", C.c_int64),
                ("page_tokens", C.c_int32), ("n_pages", C.c_int32), ("block_table", C.c_void_p),
                ("seq_lens", C.c_void_p), ("max_pages", C.c_int32), ("max_ctx", C.c_int32)]


class SpanDesc(C.Structure):
    _fields_ = [("hidden_size", C.c_int32), ("intermediate_size", C.c_int32), ("n_heads", C.c_int32),
                ("n_kv_heads", C.c_int32), ("head_dim", C.c_int32), ("n_layers", C.c_int32),
                ("dtype", C.c_int32), ("eps_input", C.c_float), ("eps_post", C.c_float),
                ("post_prescale", C.c_float)]


class LayerWeights(C.Structure):
    _fields_ = [("wqkv", C.c_void_p), ("wo", C.c_void_p), ("wgu", C.c_void_p), ("wdown", C.c_void_p),
                ("g_in", C.c_void_p), ("g_post", C.c_void_p)]


class LayerWeightsW8(C.Structure):
    _fields_ = [("wqkv", C.c_void_p), ("sqkv", C.c_void_p), ("wo", C.c_void_p), ("so", C.c_void_p),
                ("wgu", C.c_void_p), ("sgu", C.c_void_p), ("wdown", C.c_void_p), ("sdown", C.c_void_p),
                ("g_in", C.c_void_p), ("g_post", C.c_void_p)]


class ActI8(C.Structure):
    _fields_ = [("q", C.c_void_p), ("scale", C.c_void_p), ("ksum", C.c_void_p), ("q2", C.c_void_p),
                ("scale2", C.c_void_p), ("ksum2", C.c_void_p)]


class AttnMask(C.Structure):
    _fields_ = [("key_limit", C.c_void_p), ("key_valid", C.c_void_p), ("key_stride", C.c_int32)]


class StageIo(C.Structure):
    _fields_ = [("hidden_in", C.c_void_p), ("token_ids", C.c_void_p), ("embed_table", C.c_void_p),
                ("hidden_out", C.c_void_p), ("final_norm", C.c_void_p), ("lm_head", C.c_void_p),
                ("logits", C.c_void_p), ("tokens_out", C.c_void_p), ("vocab", C.c_int32), ("eps_final", C.c_float),
                ("wait_flag", C.c_void_p), ("signal_flag", C.c_void_p), ("state", C.c_void_p),
                ("wait_lag", C.c_int32), ("chain", C.c_int32), ("timeout_ms", C.c_int32)]


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found. Build it with `python -c 'import __graft_entry__ as g; g.build()'` or "
            f"`{os.path.join(_HERE, 'csrc', 'build.sh')}` (needs nvcc; targets sm_100a). "
            "There is no CPU or PyTorch fallback for the span forward.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, f32, sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t
    sig = {
        "dli_version": (C.c_int, []),
        "dli_last_error": (C.c_char_p, []),
        "dli_last_launch_count": (C.c_int, []),
        "dli_rmsnorm": (C.c_int, [vp, i32, vp, vp, vp, i32, i32, f32, f32, vp]),
        "dli_gemm": (C.c_int, [vp, i64, vp, vp, i64, i32, i32, i32, i32, vp]),
        "dli_rope_kv_append": (C.c_int, [vp, i32, vp, vp, vp, C.POINTER(KvDesc), i32, i32, i32, i32, i32, i32, vp]),
        "dli_decode_attention_scratch_bytes": (sz, [i32, i32, i32, i32]),
        "dli_decode_attention": (C.c_int, [vp, vp, C.POINTER(KvDesc), i32, i32, i32, i32, i32, vp, sz, vp]),
        "dli_decode_attention_static": (C.c_int, [vp, vp, C.POINTER(KvDesc), i32, i32, i32, i32, i32, vp]),
        "dli_prefill_attention": (C.c_int, [vp, vp, C.POINTER(KvDesc), i32, i32, i32, i32, i32, i32, i32, vp]),
        "dli_kv_evict": (C.c_int, [C.POINTER(KvDesc), i32, i32, i32, i32, i32, i32, i32, i32, vp, vp, vp]),
        "dli_silu_mul": (C.c_int, [vp, vp, i32, i32, i32, vp]),
        "dli_cast_f32_bf16": (C.c_int, [vp, vp, i64, vp]),
        "dli_span_create": (C.c_int, [C.POINTER(SpanDesc), C.POINTER(vp)]),
        "dli_span_destroy": (C.c_int, [vp]),
        "dli_span_bind_weights": (C.c_int, [vp, i32, C.POINTER(LayerWeights)]),
        "dli_span_workspace_bytes": (sz, [vp, i32, i32, i32]),
        "dli_span_profile": (C.c_int, [vp, i32]),
        "dli_span_profile_read": (C.c_int, [vp, C.POINTER(C.c_float), C.POINTER(C.c_int)]),
        "dli_span_forward": (C.c_int, [vp, C.POINTER(KvDesc), i32, i32, i32, vp, vp, vp, vp, vp, vp, sz, vp]),
        "dli_span_decode_step": (C.c_int, [vp, C.POINTER(KvDesc), i32, vp, vp, vp, vp, vp, i32, vp, sz, vp]),
        "dli_stage_decode_step": (C.c_int, [vp, C.POINTER(KvDesc), i32, C.POINTER(StageIo), vp, vp, vp, i32, vp, sz, vp]),
        "dli_flag_wait": (C.c_int, [vp, C.c_uint32, vp, i32, vp]),
        "dli_quantize_rows_w8": (C.c_int, [vp, i64, i64, vp, vp, vp]),
        "dli_dequantize_rows_w8": (C.c_int, [vp, vp, i64, i64, vp, vp]),
        "dli_gemm_w8": (C.c_int, [vp, i64, vp, vp, vp, i64, i32, i32, i32, i32, vp]),
        "dli_span_bind_weights_w8": (C.c_int, [vp, i32, C.POINTER(LayerWeightsW8)]),
        "dli_span_dequant_scratch_bytes": (sz, [vp]),
        "dli_span_set_dequant_scratch": (C.c_int, [vp, vp, sz]),
        "dli_act_i8_bytes": (sz, [i32, i32, i32]),
        "dli_act_i8_carve": (C.c_int, [vp, sz, i32, i32, i32, C.POINTER(ActI8)]),
        "dli_quantize_act_i8": (C.c_int, [vp, i64, i32, i32, f32, C.POINTER(ActI8), vp]),
        "dli_gemm_i8": (C.c_int, [C.POINTER(ActI8), vp, vp, vp, i64, i32, i32, i32, i32, vp]),
        "dli_span_act_i8_scratch_bytes": (sz, [vp]),
        "dli_span_set_act_i8": (C.c_int, [vp, vp, sz, f32]),
        "dli_span_forward_masked": (C.c_int, [vp, C.POINTER(KvDesc), i32, i32, i32, C.POINTER(AttnMask), vp, vp, vp, vp, vp, vp, sz, vp]),
        "dli_prefill_attention_masked": (C.c_int, [vp, vp, C.POINTER(KvDesc), i32, i32, i32, i32, i32, i32, i32, C.POINTER(AttnMask), vp]),
        "dli_decode_attention_masked": (C.c_int, [vp, vp, C.POINTER(KvDesc), i32, i32, i32, i32, i32, C.POINTER(AttnMask), vp, sz, vp]),
        "dli_lm_head_argmax": (C.c_int, [vp, i32, vp, f32, vp, i32, i32, i32, vp, vp, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


lib = _load()


PROF_CLASSES = ("rmsnorm", "gemm_qkv", "rope_append", "attention", "gemm_o", "gemm_gate_up", "silu_mul",
                "gemm_down", "cast")


def check(rc: int) -> None:
    if rc != DLI_OK:
        raise DliError(f"libdli_b200 error {rc}: {lib.dli_last_error().decode(errors='replace')}")


def current_stream(device=None) -> int:
    import torch

    return torch.cuda.current_stream(device).cuda_stream
