"""ctypes binding of libdllm_b200.so (C ABI: include/dllm_b200.h).

PyTorch is plumbing here (device memory, streams); every compute call goes through this C ABI into the
hand-written sm_100a kernels.  There is no fallback: if the library is missing or cannot be loaded,
importing the binding raises, and without a CUDA device every compute entry point fails.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB_PATH = os.environ.get("DLLM_B200_LIB", os.path.join(_PKG_ROOT, "lib", "libdllm_b200.so"))

ARCH_IDS = {"gpt2": 0, "llama": 1, "qwen2": 2}
DTYPE_BF16, DTYPE_F32 = 0, 1


class ShardConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("arch", "dtype", "hidden", "n_heads", "n_kv_heads", "head_dim", "ffn",
                                         "vocab", "n_layers", "has_embed", "has_head")] + \
               [("norm_eps", C.c_float)] + \
               [(n, C.c_int32) for n in ("max_rows", "max_seqs", "max_context", "page_size", "num_pages",
                                         "first_layer_mlp_only", "last_layer_attn_only")]


LAYER_FIELDS = ("norm1_w", "norm1_b", "wq", "bq", "wk", "bk", "wv", "bv", "wo", "bo", "norm2_w", "norm2_b",
                "w_gate", "w_up", "b_up", "w_down", "b_down")
IO_FIELDS = ("wte", "wpe", "final_norm_w", "final_norm_b", "lm_head", "rope_cos", "rope_sin")


class LayerWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in LAYER_FIELDS]


class IoWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in IO_FIELDS]


# name -> (restype, argtypes); must list every DLLM_API symbol of include/dllm_b200.h
_vp, _i32, _sz = C.c_void_p, C.c_int32, C.c_size_t
SIGNATURES = {
    "dllm_last_error": (C.c_char_p, []),
    "dllm_abi_version": (C.c_int, []),
    "dllm_device_count": (C.c_int, []),
    "dllm_kernel_launch_count": (C.c_ulonglong, []),
    "dllm_shard_create": (C.c_int, [C.POINTER(ShardConfig), C.POINTER(LayerWeights), C.POINTER(IoWeights),
                                    C.POINTER(_vp)]),
    "dllm_shard_destroy": (C.c_int, [_vp]),
    "dllm_shard_device_bytes": (_sz, [_vp]),
    "dllm_shard_kv_reserve": (C.c_int, [_vp, _i32, _i32, _vp]),
    "dllm_shard_kv_release": (C.c_int, [_vp, _i32]),
    "dllm_shard_kv_free_pages": (C.c_int, [_vp]),
    "dllm_shard_forward": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _vp, _vp, _vp]),
    "dllm_shard_decode_loop": (C.c_int, [_vp, _i32, _i32, _vp, _vp]),
    "dllm_shard_decode_pipelined_check": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32]),
    "dllm_shard_pipeline_status": (C.c_int, [_vp, _vp]),
    "dllm_handoff_peer_copy": (C.c_int, [_vp, C.c_int, _vp, C.c_int, _sz, _vp]),
    "dllm_mailbox_create": (C.c_int, [_i32, _i32, _i32, _i32, C.POINTER(C.c_void_p)]),
    "dllm_mailbox_export": (C.c_int, [_vp, _vp]),
    "dllm_mailbox_import": (C.c_int, [_i32, _vp, _i32, _i32, _i32, C.POINTER(C.c_void_p)]),
    "dllm_mailbox_view": (C.c_int, [_vp, _i32, C.POINTER(C.c_void_p)]),
    "dllm_mailbox_post_tokens": (C.c_int, [_vp, _i32, _vp, C.c_uint32, _vp]),
    "dllm_mailbox_read_tokens": (C.c_int, [_vp, _i32, _vp, _vp, _vp]),
    "dllm_mailbox_put_rows": (C.c_int, [_vp, _i32, _vp, _i32, _vp]),
    "dllm_mailbox_get_rows": (C.c_int, [_vp, _i32, _vp, _i32, _vp]),
    "dllm_mailbox_destroy": (C.c_int, [_vp]),
    "dllm_shard_decode_pipelined": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, C.c_uint32, _vp, _vp, _vp, _vp]),
    "dllm_op_gemv": (C.c_int, [_i32, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "dllm_op_gemm": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "dllm_op_decode_attention_scratch_bytes": (_sz, [_i32, _i32, _i32, _i32]),
    "dllm_op_decode_attention": (C.c_int, [_i32, _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _i32, _i32, _i32, _i32,
                                           _i32, _vp, _vp, _vp]),
    "dllm_op_decode_attention_static": (C.c_int, [_i32, _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _i32, _i32, _i32, _i32,
                                                  _i32, _vp, _vp]),
    "dllm_op_legacy_matmul": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp]),
}

_lib = None
_lock = threading.Lock()


class DllmError(RuntimeError):
    """A C-ABI call returned non-zero; the message is dllm_last_error()."""


def lib() -> C.CDLL:
    """Load the shared library once.  Raises (never falls back) when it is absent."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise DllmError(
                        f"{LIB_PATH} not found: build it with `python distributed-llms_b200/build.py` "
                        "(there is no CPU or PyTorch fallback for the shard forward pass)")
                l = C.CDLL(LIB_PATH)
                for name, (res, args) in SIGNATURES.items():
                    fn = getattr(l, name)
                    fn.restype, fn.argtypes = res, args
                _lib = l
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise DllmError(lib().dllm_last_error().decode(errors="replace"))


def device_count() -> int:
    return int(lib().dllm_device_count())


def kernel_launch_count() -> int:
    return int(lib().dllm_kernel_launch_count())
