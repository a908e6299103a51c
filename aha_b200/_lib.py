"""ctypes binding of libaha_b200.so (the C-ABI in include/aha_b200.h).  There is no fallback: if the CUDA library is
missing or a call fails, this module raises."""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_double, c_float, c_int, c_int32, c_longlong, c_ulonglong, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libaha_b200.so")

POLICY_SINK, POLICY_SLIDING, POLICY_STATIC = 0, 1, 2
MASK_SDPA, MASK_FA2 = 0, 1
LOGITS_NONE, LOGITS_LAST, LOGITS_ALL = 0, 1, 2
POOL_MODES = {"bilinear": 0, "average": 1, "max": 2, "adaptive": 3}
ACT_NONE, ACT_GELU_TANH, ACT_GELU_ERF, ACT_SILU_MUL = 0, 1, 2, 3


class ModelDesc(Structure):
    _fields_ = [(n, c_int) for n in ("hidden", "inter", "layers", "heads", "kv_heads", "head_dim", "vocab")] + \
               [("rope_theta", c_float), ("rms_eps", c_float)] + \
               [(n, c_int) for n in ("v_hidden", "v_inter", "v_layers", "v_heads", "v_patch", "v_image")] + \
               [("v_eps", c_float)] + \
               [(n, c_int) for n in ("pool_stride", "pool_mode", "max_tokens", "max_frames")]


class WeightEntry(Structure):
    _fields_ = [("name", c_char_p), ("ptr", c_void_p)]


# every symbol include/aha_b200.h declares: name -> (restype, argtypes)
P = c_void_p
SIGNATURES = {
    "aha_last_error": (c_char_p, []),
    "aha_launch_count": (c_ulonglong, []),
    "aha_set_sm_sharing": (c_int, [c_int]),
    "aha_debug_attention_timeline": (c_int, [POINTER(c_ulonglong)]),
    "aha_debug_attention_persistent_timeline": (c_int, [POINTER(c_ulonglong)]),
    "aha_attention_persistent_launch_count": (c_ulonglong, []),
    "aha_debug_gemm_timeline": (c_int, [POINTER(c_int), POINTER(c_int), POINTER(c_int), POINTER(c_ulonglong)]),
    "aha_debug_gemm_tails": (c_int, [c_int, POINTER(c_ulonglong)]),
    "aha_profile_gemm": (c_int, [c_int]),
    "aha_profile_gemm_read": (c_int, [c_int, POINTER(c_double), POINTER(c_longlong), POINTER(c_double), POINTER(c_double)]),
    "aha_model_create": (c_int, [POINTER(ModelDesc), POINTER(WeightEntry), c_int, c_int, POINTER(P)]),
    "aha_model_destroy": (c_int, [P]),
    "aha_cache_create": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, POINTER(P)]),
    "aha_cache_reset": (c_int, [P]),
    "aha_cache_destroy": (c_int, [P]),
    "aha_cache_update": (c_int, [P, c_int, P, P, c_longlong, c_longlong, c_longlong, c_longlong, P, P, c_int, P]),
    "aha_cache_seq_length": (c_int, [P, c_int, POINTER(c_int)]),
    "aha_cache_seen_tokens": (c_int, [P, POINTER(c_longlong)]),
    "aha_cache_export": (c_int, [P, c_int, P, P, P, P]),
    "aha_vision_embed": (c_int, [P, P, P, c_int, P, P]),
    "aha_assemble_chunk": (c_int, [P, P, c_int, P, c_int, P, P]),
    "aha_prefill_chunk": (c_int, [P, POINTER(P), c_int, P, c_int, c_int, c_int, P, P, P, P, P, P, P]),
    "aha_find_ticks": (c_int, [P, c_int, c_int, c_double, c_int, P, P, c_int, P, P, P, P]),
    "aha_greedy_pick": (c_int, [P, c_int, P, c_int, c_float, P, P, P]),
    "aha_fuse_scores": (c_int, [P, P, P, c_longlong, c_double, c_double, c_double, c_double, P, P]),
    "aha_knapsack_words": (c_int, [c_int]),
    "aha_knapsack_select": (c_int, [P, P, P, c_int, c_int, c_int, c_double, c_double, c_double, P, P, P]),
    "aha_highlight_mask": (c_int, [P, P, c_int, c_int, c_int, c_int, P, P]),
    "aha_gemm_bf16": (c_int, [P, c_int, c_int, P, c_int, c_int, c_int, P, c_int, c_int, c_int, c_int, P, P, c_int, c_int,
                              c_int, c_int, c_int, P]),
    "aha_rmsnorm": (c_int, [P, P, P, c_int, c_int, c_float, P]),
    "aha_layernorm": (c_int, [P, P, P, P, c_int, c_int, c_float, P]),
    "aha_preprocess": (c_int, [P, P, c_longlong, P]),
    "aha_ingest_frames": (c_int, [P, c_int, c_int, c_int, c_int, P, c_int, P]),
    "aha_pool": (c_int, [P, P, c_int, c_int, c_int, c_int, c_int, P]),
    "aha_rope_table": (c_int, [c_int, c_int, c_int, c_float, P, P, P]),
    "aha_attention": (c_int, [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_float, P]),
    "aha_attention_noalloc": (c_int, [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_float, P]),
}

_lib = None


def load():
    """Loads the shared library (once) and types every entry point.  Raises if it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU / PyTorch fallback for the aha_b200 hot path)")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int):
    """Maps the C-ABI status codes back to the exception the reference would raise."""
    if rc == 0:
        return
    msg = (load().aha_last_error() or b"").decode()
    if rc == -1:
        raise ValueError(msg.replace("ValueError: ", "", 1))
    raise RuntimeError(msg.replace("RuntimeError: ", "", 1))


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count() -> int:
    return int(load().aha_launch_count())
