"""ctypes binding of libacai_b200.so (the C ABI declared in include/acai_b200.h).

The library is built in-tree by `__graft_entry__.build()` (or `make -C acai_omr_b200/csrc`).
There is no fallback: if the shared object is missing or no sm_100 device is present, every
op raises.  This module is also the reference-side binding shown in INTEGRATION.md.
"""
from __future__ import annotations

import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ACAI_LIB_PATH") or os.path.join(_HERE, "libacai_b200.so")     # (override: kernel A/B experiments)
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "acai_b200.h")

_lib = None


class AcaiError(RuntimeError):
    pass


class ImageDesc(C.Structure):
    """acai_image_desc"""
    _fields_ = [("pixels", C.c_void_p), ("height", C.c_int32), ("width", C.c_int32),
                ("hp", C.c_int32), ("wp", C.c_int32)]


class ResizeDesc(C.Structure):
    """acai_resize_desc"""
    _fields_ = [("src", C.c_void_p), ("src_dtype", C.c_int32), ("src_h", C.c_int32), ("src_w", C.c_int32),
                ("src_stride", C.c_int64), ("tmp", C.c_void_p), ("dst", C.c_void_p), ("dst_h", C.c_int32), ("dst_w", C.c_int32),
                ("crop_top", C.c_int32), ("crop_left", C.c_int32), ("out_h", C.c_int32), ("out_w", C.c_int32),
                ("dst_stride", C.c_int64)]


_p, _i32, _i64, _f32, _sz, _u64 = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t, C.c_uint64

# name -> argtypes, in header order
SIGNATURES = {
    "acai_version": [],
    "acai_last_error": [],
    "acai_device_ok": [],
    "acai_patch_embed_workspace_bytes": [_i32, _i32, _i32, _i32],
    "acai_patch_embed_fwd": [_p, _i32, _p, _p, _i32, _i32, _i32, _p, _p, _p, _i32, _i32, _i32, _p, _i32, _i32, _p, _sz, _p],
    "acai_linear_fwd": [_p, _i32, _i64, _p, _i32, _i64, _p, _p, _i32, _i64, _p, _i32, _i64, _i32, _i32, _i32, _i32, _p],
    "acai_linear_layernorm": [_p, _i64, _p, _i64, _p, _p, _i32, _i64, _p, _p, _f32, _p, _i64, _p, _i64, _i32, _i32, _i32, _p],
    "acai_linear_layernorm_resident_clusters": [_i32],
    "acai_layernorm_fwd": [_p, _i32, _p, _i32, _p, _p, _f32, _p, _i32, _p, _i32, _i32, _i32, _p],
    "acai_attn_varlen_fwd": [_p, _p, _p, _i32, _i64, _i64, _i64, _p, _p, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _f32, _p, _i32,
                             _i64, _p, _f32, _u64, _p],
    "acai_attn_dropout_mask": [_f32, _u64, _i32, _i32, _i32, _p, _p],
    "acai_attn_varlen_bwd": [_p, _p, _p, _i32, _i64, _i64, _i64, _p, _i64, _p, _i64, _p, _p, _p, _i32, _i32, _i32, _i32, _i32, _i32,
                             _i32, _i32, _f32, _p, _p, _p, _i32, _i64, _i64, _i64, _p, _f32, _u64, _p],
    "acai_transpose": [_p, _i32, _i64, _i32, _i32, _p, _i64, _p],
    "acai_colsum_workspace_bytes": [_i32, _i32],
    "acai_colsum": [_p, _i32, _i64, _i32, _i32, _p, _i32, _p, _sz, _p],
    "acai_colsum_static_max_workspace_bytes": [],
    "acai_colsum_static": [_p, _i32, _i64, _i32, _i32, _p, _i32, _p],
    "acai_gelu_fwd": [_p, _i32, _p, _i32, _i64, _f32, _u64, _p],
    "acai_gelu_bwd": [_p, _i32, _p, _i32, _p, _i32, _i64, _f32, _u64, _p],
    "acai_linear_dropout": [_p, _i64, _p, _i64, _p, _p, _i32, _i64, _p, _i64, _i32, _i32, _i32, _f32, _u64, _p],
    "acai_layernorm_bwd_workspace_bytes": [_i32, _i32],
    "acai_layernorm_bwd": [_p, _i32, _p, _i32, _p, _i32, _p, _f32, _p, _i32, _p, _p, _i32, _i32, _i32, _p, _sz, _p],
    "acai_ce_loss": [_p, _i32, _i64, _p, _i32, _i32, _i32, _p, _p, _p, _p, _i32, _i64, _p, _f32, _p],
    "acai_scatter_add_rows": [_p, _i32, _i64, _p, _i32, _i32, _p, _f32, _p],
    "acai_gather_rows": [_p, _i32, _p, _i32, _i32, _p, _i32, _p],
    "acai_mae_loss_bwd": [_p, _i32, _p, _p, _i32, _i32, _p, _p, _i32, _p],
    "acai_dropout": [_p, _i32, _p, _i32, _i64, _f32, _u64, _p, _i32, _p],
    "acai_dropout_mask": [_i64, _f32, _u64, _p, _p],
    "acai_gumbel_softmax_fwd": [_p, _i32, _i64, _p, _i32, _i32, _f32, _i32, _p, _p, _i32, _p],
    "acai_gumbel_softmax_bwd": [_p, _i32, _p, _i32, _i32, _f32, _p, _i32, _p],
    "acai_mix_rows": [_p, _i32, _p, _p, _i32, _p, _i32, _i32, _i32, _p, _i32, _p],
    "acai_dynamic_resize": [_p, _i32, _i32, _i32, _i32, _i32, _p],
    "acai_linear_splitk_workspace_bytes": [_i32, _i32, _i32],
    "acai_linear_splitk": [_p, _i64, _p, _i64, _p, _p, _i32, _i64, _p, _p, _f32, _p, _i32, _i64, _p, _i32, _i64, _i32, _i32, _i32,
                           _i32, _p, _sz, _p],
    "acai_linear_gelu_train": [_p, _i64, _p, _i64, _p, _p, _i64, _p, _i64, _i32, _i32, _i32, _f32, _u64, _p],
    "acai_linear_gelu_bwd": [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _i32, _i32, _i32, _f32, _u64, _p],
    "acai_linear_dgrad": [_p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _i64, _i32, _i32, _i32, _f32, _u64, _p],
    "acai_wgrad_accumulate": [_p, _i64, _p, _i64, _i32, _p, _i64, _i32, _i32, _i32, _p],
    "acai_wgrad_bias_accumulate": [_p, _i64, _p, _i64, _i32, _p, _i64, _p, _i32, _i32, _i32, _p],
    "acai_adamw_step": [_p, _p, _p, _p, _p, _i64, _f32, _f32, _f32, _f32, _f32, _i32, _f32, _p],
    "acai_decode_attn": [_p, _i32, _i64, _p, _p, _i32, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p, _i64, _f32, _p,
                         _i32, _i64, _p, _sz, _p],
    "acai_kv_append": [_p, _p, _i32, _i64, _i64, _i64, _i32, _p, _p, _i32, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p],
    "acai_decode_workspace_bytes": [_i32, _i32, _i32],
    "acai_decode_attn_paged": [_p, _i64, _p, _i64, _p, _i32, _p, _i32, _p, _p, _p, _i32, _i32, _i32, _i32, _i32, _f32, _p,
                               _i32, _i64, _p, _sz, _i64, _p],
    "acai_decode_embed": [_p, _i64, _p, _i32, _p, _p, _p, _i32, _p, _i32, _i32, _p],
    "acai_greedy_update": [_p, _i32, _i64, _i32, _i32, _p, _p, _i64, _p, _p, _i32, _p],
    "acai_topk_sample_update": [_p, _i32, _i64, _i32, _i32, _i32, _f32, _p, _i64, _p, _p, _p, _i64, _p, _p, _i32, _p],
    "acai_advance_step": [_p, _p],
    "acai_embed_tokens": [_p, _i32, _p, _p, _p, _i32, _i32, _p, _i32, _i32, _p],
    "acai_argmax_logprob": [_p, _i32, _i64, _i32, _i32, _p, _p, _p],
    "acai_unpack_padded": [_p, _i32, _p, _i32, _i32, _i32, _p, _p, _i32, _p],
    "acai_patchify": [_p, _i32, _p, _i32, _i32, _i32, _p, _i32, _p],
    "acai_mae_restore": [_p, _i32, _p, _p, _p, _p, _i32, _i32, _p, _p, _i32, _p, _i32, _i32, _p],
    "acai_mae_loss_fwd": [_p, _i32, _p, _p, _i32, _i32, _p, _p, _p],
    "acai_pack_padded": [_p, _i32, _p, _i32, _i32, _i32, _p, _i32, _p],
}


def header_symbols():
    """Every function name include/acai_b200.h declares."""
    with open(HEADER_PATH) as f:
        text = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(acai_[a-z0-9_]+)\s*\(", text)))


def load():
    """Load the shared library (once).  Raises AcaiError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise AcaiError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  acai_omr_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = (C.c_char_p if name == "acai_last_error" else
                      C.c_size_t if name.endswith("_workspace_bytes") else C.c_int)
    _lib = lib
    return lib


def check(status, what):
    if status != 0:
        msg = load().acai_last_error()
        raise AcaiError(f"{what} failed ({status}): {msg.decode() if msg else '?'}")


_device_checked = False


def require_device():
    """Fail loudly unless an sm_100 device is usable."""
    global _device_checked
    if _device_checked:
        return
    lib = load()
    if not lib.acai_device_ok():
        raise AcaiError("acai_omr_b200 needs an NVIDIA B200 (sm_100) device: "
                        + lib.acai_last_error().decode() + "; there is no CPU fallback")
    _device_checked = True
