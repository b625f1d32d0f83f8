"""ctypes binding of libb200gguf.so (the C ABI in include/b200_gguf.h).

The counterpart of llama-cpp-python's `llama_cpp/llama_cpp.py` ctypes layer that the
reference imports through `from llama_cpp import Llama` (/root/reference/api.py:7).
There is no CPU fallback: if the library is missing or fails to load this raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libb200gguf.so")

B2_FLAG_NO_GRAPH, B2_FLAG_NO_PDL, B2_FLAG_ACT8 = 1, 2, 4


class B2Config(ctypes.Structure):
    _fields_ = [(n, c_int32) for n in ("n_layer", "n_embd", "n_head", "n_head_kv", "head_dim", "n_ff", "n_vocab")] + [
        ("rms_eps", c_float), ("rope_freq_base", c_float)] + [
        (n, c_int32) for n in ("n_ctx", "max_seqs", "kv_block_tokens", "kv_num_blocks", "device", "flags")]


class B2Sampling(ctypes.Structure):
    _fields_ = [("temperature", c_float), ("top_p", c_float), ("min_p", c_float), ("repeat_penalty", c_float),
                ("frequency_penalty", c_float), ("presence_penalty", c_float), ("top_k", c_int32),
                ("penalty_last_n", c_int32)]


# name -> (restype, argtypes); every symbol include/b200_gguf.h declares
SIGNATURES = {
    "b2_last_error": (c_char_p, []),
    "b2_version": (c_int, []),
    "b2_device_count": (c_int, [POINTER(c_int)]),
    "b2_engine_create": (c_int, [POINTER(B2Config), POINTER(c_void_p)]),
    "b2_engine_destroy": (c_int, [c_void_p]),
    "b2_upload_tensor": (c_int, [c_void_p, c_char_p, c_int, c_int64, c_int64, c_void_p, c_int64]),
    "b2_tp_unique_id": (c_int, [c_void_p]),
    "b2_engine_tp_init": (c_int, [c_void_p, c_int, c_int, c_void_p]),
    "b2_engine_finalize": (c_int, [c_void_p]),
    "b2_engine_device_bytes": (c_int, [c_void_p, POINTER(c_int64), POINTER(c_int64)]),
    "b2_seq_set_blocks": (c_int, [c_void_p, c_int, POINTER(c_int32), c_int]),
    "b2_seq_begin": (c_int, [c_void_p, c_int, c_int, POINTER(c_int32), c_int, POINTER(B2Sampling), POINTER(c_double), c_int]),
    "b2_prefill": (c_int, [c_void_p, c_int, POINTER(c_int32), c_int]),
    "b2_sample": (c_int, [c_void_p, c_int, POINTER(c_int32)]),
    "b2_decode": (c_int, [c_void_p, POINTER(c_int32), c_int, POINTER(c_int32)]),
    "b2_decode_n": (c_int, [c_void_p, c_int, c_int, POINTER(c_int32)]),
    "b2_seq_rewind": (c_int, [c_void_p, c_int, c_int, c_int32, POINTER(c_int32), c_int]),
    "b2_seq_set_token": (c_int, [c_void_p, c_int, c_int32]),
    "b2_forward_pending": (c_int, [c_void_p, c_int]),
    "b2_get_logits": (c_int, [c_void_p, c_int, c_void_p]),
    "b2_synchronize": (c_int, [c_void_p]),
    "b2_timer_start": (c_int, [c_void_p]),
    "b2_timer_stop": (c_int, [c_void_p, POINTER(c_float)]),
    "b2_kernel_launch_count": (c_int, [c_void_p, POINTER(c_int64)]),
    "b2_engine_mode": (c_int, [c_void_p, POINTER(c_int), POINTER(c_int), POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "b2_engine_tp_mode": (c_int, [c_void_p, POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "b2_profile_decode_step": (c_int, [c_void_p, c_int, POINTER(c_float), POINTER(c_int32), POINTER(c_int64)]),
    "b2_bench_matvec": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, POINTER(c_float)]),
    "b2_op_dequantize": (c_int, [c_int, c_int, c_void_p, c_int64, c_void_p]),
    "b2_op_matvec": (c_int, [c_int, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_int, c_int]),
    "b2_op_gate_up": (c_int, [c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p, c_float, c_void_p, c_int]),
    "b2_op_gemm": (c_int, [c_int, c_int, c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int]),
    "b2_op_rmsnorm": (c_int, [c_int, c_void_p, c_void_p, c_float, c_int, c_int, c_void_p]),
    "b2_op_attention_decode": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int, c_int, c_void_p]),
    "b2_attention_decode_static": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int,
                                           c_int, c_int, c_void_p, c_void_p]),
    "b2_op_sample": (c_int, [c_int, c_void_p, c_int, POINTER(B2Sampling), c_void_p, c_int, c_double, POINTER(c_int32)]),
}

_lib = None


def load_library(path: str | None = None) -> ctypes.CDLL:
    """Load libb200gguf.so (building nothing: see build.py). Raises if it is not there."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("B200_GGUF_LIB") or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(
            f"{p} not found: the B200 engine has no CPU fallback. Build it with "
            f"`python {os.path.join(os.path.dirname(_HERE), 'build.py')}` (needs nvcc).")
    lib = ctypes.CDLL(p)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


class B2Error(RuntimeError):
    pass


def check(rc: int, what: str = "") -> None:
    """Translate a negative return code into a Python exception (RuntimeError family)."""
    if rc == 0:
        return
    msg = load_library().b2_last_error()
    text = msg.decode("utf-8", "replace") if msg else ""
    raise B2Error(f"{what or 'b200 engine call'} failed (code {rc}): {text}")
