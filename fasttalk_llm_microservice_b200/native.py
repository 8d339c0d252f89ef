"""ctypes binding of libfasttalk_b200.so (the C ABI declared in include/fasttalk_b200.h).

There is no Python / CPU fallback: if the library is missing or a call fails, this
module raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libfasttalk_b200.so")


class NativeError(RuntimeError):
    pass


class EngineConfig(C.Structure):
    _fields_ = [
        ("hidden", C.c_int), ("ffn", C.c_int), ("layers", C.c_int), ("n_q_heads", C.c_int),
        ("n_kv_heads", C.c_int), ("head_dim", C.c_int), ("vocab", C.c_int),
        ("rms_eps", C.c_float),
        ("max_batch", C.c_int), ("max_batched_tokens", C.c_int), ("num_pages", C.c_int),
        ("max_context", C.c_int), ("max_sessions", C.c_int),
        ("n_eos", C.c_int), ("eos_ids", C.c_int * 4),
        ("use_pdl", C.c_int), ("debug_logit_rows", C.c_int),
        ("tp_size", C.c_int), ("tp_rank", C.c_int),
        ("decode_prefill_cap", C.c_int), ("prefill_planes", C.c_int),
    ]


class Weights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "embed", "lm_head", "final_norm", "attn_norm", "wqkv", "wo", "mlp_norm", "wgu", "wd",
        "rope_cos", "rope_sin")]


class EngineStats(C.Structure):
    _fields_ = [
        ("steps", C.c_longlong), ("decode_steps", C.c_longlong), ("prefill_tokens", C.c_longlong),
        ("decode_tokens", C.c_longlong), ("kernel_launches", C.c_longlong),
        ("prefix_hit_tokens", C.c_longlong),
        ("h2d_bytes", C.c_longlong), ("d2h_bytes", C.c_longlong), ("decode_ctx_tokens", C.c_longlong),
        ("attn_decode_launches", C.c_longlong), ("attn_decode_kv_tokens", C.c_longlong),
        ("attn_decode_ms", C.c_double),
        ("active_requests", C.c_int), ("waiting_requests", C.c_int), ("live_sessions", C.c_int),
        ("free_pages", C.c_int), ("total_pages", C.c_int), ("last_batch", C.c_int),
        ("last_prefill_tokens", C.c_int), ("last_step_ms", C.c_double),
    ]


class GemmEpilogue(C.Structure):
    """ft_gemm_epilogue (include/fasttalk_b200.h)."""
    _fields_ = [
        ("mode", C.c_int), ("out", C.c_void_p), ("out_lo", C.c_void_p), ("resid", C.c_void_p),
        ("gamma", C.c_void_p), ("ssq_out", C.c_void_p), ("ssq_in", C.c_void_p),
        ("ssq_parts", C.c_int), ("ssq_stride", C.c_int), ("inv_h", C.c_float), ("eps", C.c_float),
        ("rope_cos", C.c_void_p), ("rope_sin", C.c_void_p), ("positions", C.c_void_p),
        ("seq_rows", C.c_void_p), ("page_table", C.c_void_p), ("max_pages", C.c_int),
        ("q_out", C.c_void_p), ("kv_layer", C.c_void_p),
        ("nq", C.c_int), ("nkv", C.c_int), ("head_dim", C.c_int),
        ("scratch", C.c_void_p), ("sync", C.c_void_p),
        ("prefetch", C.c_void_p), ("prefetch_bytes", C.c_longlong),
        ("tp_peer_xchg", C.c_void_p), ("tp_xchg", C.c_void_p), ("tp_peer_flags", C.c_void_p),
        ("tp_flags", C.c_void_p), ("tp_epoch", C.c_uint), ("tp_peer_out", C.c_void_p),
        ("out_stride", C.c_int), ("out_col0", C.c_int), ("sync_mode", C.c_int), ("sync_slots", C.c_int),
        ("out_f16", C.c_void_p), ("x_planes", C.c_int),
    ]


EPI_PARTIAL, EPI_BF16, EPI_SWIGLU, EPI_RESID, EPI_QKV_ROPE, EPI_LOGITS = range(6)

# name -> (restype, argtypes); every symbol include/fasttalk_b200.h declares
_VP, _I, _LL, _F, _ULL = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong
SIGNATURES = {
    "ft_last_error": (C.c_char_p, []),
    "ft_engine_create": (_VP, [C.POINTER(EngineConfig), C.POINTER(Weights)]),
    "ft_engine_destroy": (None, [_VP]),
    "ft_engine_submit": (_LL, [_VP, _ULL, _VP, _I, _F, _I, _F, _I, _ULL, _I]),
    "ft_engine_step": (_I, [_VP]),
    "ft_engine_start": (_I, [_VP]),
    "ft_engine_stop": (_I, [_VP]),
    "ft_engine_wait": (_I, [_VP, _I]),
    "ft_engine_drain": (_I, [_VP, _VP, _VP, _VP, _I]),
    "ft_engine_cancel": (_I, [_VP, _LL]),
    "ft_engine_session_close": (_I, [_VP, _ULL]),
    "ft_engine_stats_get": (_I, [_VP, C.POINTER(EngineStats)]),
    "ft_engine_get_logits": (_I, [_VP, _LL, _VP]),
    "ft_engine_timer": (_I, [_VP, _I, _VP]),
    "ft_engine_set_profile": (_I, [_VP, _I]),
    "ft_engine_set_decode_prefill_cap": (_I, [_VP, _I]),
    "ft_engine_step_log": (_I, [_VP, _VP, _I]),
    "ft_debug_prefill_logits": (_I, [_VP, _VP, _I, _I, _VP]),
    "ft_engine_tp_export": (_I, [_VP, _VP, _VP, _VP]),
    "ft_engine_tp_attach": (_I, [_VP, _VP, _VP, _VP]),
    "ft_ipc_export": (_I, [_VP, _VP]),
    "ft_ipc_open": (_I, [_VP, _VP]),
    "ft_ipc_close": (_I, [_VP]),
    "ft_gemm_bf16": (_I, [_VP, _VP, _VP, _VP, _VP, _I, _I, _I, _I, _I, _VP]),
    "ft_gemm_pick_splits": (_I, [_I, _I, _I, _I]),
    "ft_debug_gemm_trace": (_I, [_VP]),
    "ft_debug_gemm_trace_hist": (_I, [_VP, _I]),
    "ft_gemm_scratch_bytes": (_LL, [_I, _I, _I, _I]),
    "ft_gemm_fused": (_I, [_VP, _VP, _VP, _I, _I, _I, _I, _VP, _VP]),
    "ft_embed_planes": (_I, [_VP, _VP, _VP, _VP, _VP, _VP, _VP, _I, _I, _I, _I, _VP]),
    "ft_gather_planes": (_I, [_VP, _VP, _VP, _I, _I, _VP, _I, _I, _VP, _VP, _VP, _I, _VP]),
    "ft_add_rmsnorm": (_I, [_VP, _VP, _I, _LL, _VP, _VP, _VP, _VP, _VP, _VP, _I, _I, _F, _VP]),
    "ft_rope_kv_append": (_I, [_VP, _I, _LL, _VP, _VP, _VP, _VP, _VP, _I, _VP, _VP, _I, _I, _I, _I, _VP]),
    "ft_silu_mul": (_I, [_VP, _I, _LL, _VP, _VP, _I, _I, _VP]),
    "ft_attn_decode_scratch_bytes": (_LL, [_I, _I, _I, _I, _I]),
    "ft_attn_decode_default_ctas": (_I, [_I]),
    "ft_debug_attn_trace": (_I, [_VP, _I]),
    "ft_attn_decode": (_I, [_VP, _VP, _VP, _I, _VP, _VP, _VP, _VP, _VP, _I, _I, _I, _I, _I, _VP]),
    "ft_attn_decode_noalloc": (_I, [_VP, _VP, _VP, _I, _VP, _VP, _VP, _I, _I, _I, _I, _I, _VP]),
    "ft_attn_prefill": (_I, [_VP, _VP, _VP, _I, _VP, _VP, _VP, _VP, _VP, _I, _VP, _VP, _I, _I, _I, _I, _VP]),
    "ft_sample": (_I, [_VP, _I, _I, _VP, _VP, _VP, _VP, _VP, _VP]),
}

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """dlopen the in-tree library and type every entry point; raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NativeError(
            f"{LIB_PATH} is missing: build it with `python -m fasttalk_llm_microservice_b200.build` "
            "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return (load().ft_last_error() or b"").decode("utf-8", "replace")


def check(rc: int, what: str) -> int:
    if rc < 0:
        raise NativeError(f"{what} failed ({rc}): {last_error()}")
    return rc


def ptr(t) -> int:
    """Device/host pointer of a torch tensor (None -> NULL)."""
    return 0 if t is None else t.data_ptr()


def cur_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
