"""ctypes binding of the C ABI declared in ``include/seva_b200.h``.

The shared library is built in-tree (``stable-virtual-camera_b200/lib/libseva_b200.so``) by
``make -C stable-virtual-camera_b200/csrc`` / ``__graft_entry__.build()``.  There is no fallback: if the
library is missing or the device is not sm_100 every op raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SEVA_B200_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libseva_b200.so")  # override: developer A/B builds

c_void_p, c_int, c_i64, c_float = C.c_void_p, C.c_int32, C.c_int64, C.c_float


class SevaGemm(C.Structure):
    _fields_ = [
        ("A", c_void_p), ("lda", c_i64), ("a_rows", c_i64),
        ("W", c_void_p), ("ldw", c_i64),
        ("M", c_int), ("N", c_int), ("Kc", c_int), ("ntaps", c_int), ("tap_wp", c_int),
        ("bias", c_void_p), ("rowbias", c_void_p), ("rowbias_rpi", c_int),
        ("residual", c_void_p), ("ldr", c_i64),
        ("out", c_void_p), ("ldo", c_i64),
        ("out_fp32", c_int), ("geglu", c_int), ("out_map", c_int), ("H", c_int), ("Wd", c_int),
        ("mask_border", c_int), ("Hp", c_int), ("Wp", c_int), ("bn", c_int), ("cluster", c_int),
        ("reverse", c_int),
        ("gn_partial", c_void_p), ("gn_rpi", c_int), ("gn_slots", c_int), ("ln_rowstats", c_void_p),
        ("ln_in_stats", c_void_p), ("ln_in_slots", c_int), ("ln_eps", c_float), ("ln_colsum", c_void_p),
    ]


class SevaGNApply(C.Structure):
    _fields_ = [
        ("x", c_void_p), ("ldx", c_i64), ("y", c_void_p), ("ldy", c_i64),
        ("sums", c_void_p), ("gamma", c_void_p), ("beta", c_void_p),
        ("Nimg", c_int), ("H", c_int), ("Wd", c_int), ("C", c_int), ("groups", c_int),
        ("eps", c_float), ("in_padded", c_int), ("out_padded", c_int), ("silu", c_int),
        ("dense", c_void_p), ("dense_w", c_void_p), ("dense_b", c_void_p), ("reverse", c_int),
        ("gstats", c_void_p),
    ]


class SevaGemvProblem(C.Structure):
    _fields_ = [("x", c_void_p), ("W", c_void_p), ("bias", c_void_p), ("out", c_void_p),
                ("M", c_int), ("N", c_int), ("K", c_int), ("silu_in", c_int), ("silu_out", c_int), ("pad_", c_int)]


# name -> argtypes (restype is always int unless listed in _RESTYPES)
_PROTOS = {
    "seva_check_device": [],
    "seva_version": [],
    "seva_gemm_bf16": [C.POINTER(SevaGemm), c_void_p],
    "seva_gemm_pick_bn": [c_int, c_int],
    "seva_gemm_ln_slots": [c_int, c_int],
    "seva_attn_fwd": [c_void_p] * 4 + [c_int] * 4 + [c_i64] * 8 + [c_float, c_void_p],
    "seva_attn_partial": [c_void_p] * 5 + [c_int] * 4 + [c_i64] * 6 + [c_float, c_void_p],
    "seva_attn_merge": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_i64, c_i64, c_void_p],
    "seva_timemix_attn": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_i64, c_i64, c_float, c_void_p],
    "seva_timemix_attn_kv": [c_void_p] * 4 + [c_int] * 5 + [c_i64] * 3 + [c_float, c_void_p],
    "seva_groupnorm_stats_floats": [c_int, c_int, c_int],
    "seva_groupnorm_stats": [c_void_p, c_i64, c_int, c_int, c_int, c_void_p, c_int, c_void_p],
    "seva_groupnorm_stats_static": [c_void_p, c_i64, c_int, c_int, c_int, c_void_p, c_int, c_void_p],
    "seva_groupnorm_apply": [C.POINTER(SevaGNApply), c_void_p],
    "seva_groupnorm_partial_floats": [c_int, c_int, c_int],
    "seva_groupnorm_finalize": [c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_float, c_void_p,
                                c_void_p],
    "seva_layernorm": [c_void_p, c_i64, c_void_p, c_i64, c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_void_p],
    "seva_plucker": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p],
    "seva_resize_bilinear_ac": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "seva_multiview_scale": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_float, c_void_p, c_void_p, c_int,
                             c_void_p],
    "seva_sigma_quantize": [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p],
    "seva_add_noise": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_i64, c_void_p],
    "seva_denoise_prepare": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_i64, c_void_p, c_int, c_int,
                             c_int, c_int, c_void_p],
    "seva_denoise_finish": [c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p],
    "seva_cfg_euler_step": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_i64,
                            c_void_p],
    "seva_pack_nchw": [c_void_p, c_void_p, c_i64, c_int, c_int, c_int, c_int, c_void_p],
    "seva_unpack_nchw": [c_void_p, c_i64, c_void_p, c_int, c_int, c_int, c_int, c_void_p],
    "seva_upsample2x_pad": [c_void_p, c_i64, c_void_p, c_i64, c_int, c_int, c_int, c_int, c_void_p],
    "seva_im2col3x3_s2": [c_void_p, c_i64, c_void_p, c_int, c_int, c_int, c_int, c_void_p],
    "seva_im2col3x3_s2_at": [c_void_p, c_i64, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "seva_softmax_rows": [c_void_p, c_i64, c_void_p, c_i64, c_int, c_int, c_void_p],
    "seva_cast_f32_bf16": [c_void_p, c_void_p, c_i64, c_void_p],
    "seva_timestep_embedding": [c_void_p, c_void_p, c_int, c_int, c_float, c_void_p],
    "seva_gemv_grouped": [c_void_p, c_int, c_int, c_int, c_int, c_void_p],
    "seva_gemv_rows": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "seva_emb_gather": [c_void_p, c_i64, c_int, c_void_p, c_int, c_void_p, c_int, c_void_p],
    "seva_resize_bicubic_aa": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "seva_clip_patchify": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, C.POINTER(c_float), C.POINTER(c_float),
                           c_void_p],
    "seva_resize_area": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p],
}
EXPORTED_SYMBOLS = sorted(list(_PROTOS) + ["seva_last_error"])

_lib = None


class SevaB200Error(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load the library (once).  Raises if it has not been built: there is no other code path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SevaB200Error(
            f"{LIB_PATH} not found: build it with `make -C stable-virtual-camera_b200/csrc` "
            "(or __graft_entry__.build()); seva_b200 has no CPU / PyTorch fallback"
        )
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in _PROTOS.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = C.c_int
    lib.seva_groupnorm_stats_floats.restype = C.c_int64
    lib.seva_last_error.argtypes = []
    lib.seva_last_error.restype = C.c_char_p
    _lib = lib
    return lib


def last_error() -> str:
    return load().seva_last_error().decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise SevaB200Error(f"{what} failed (rc={rc}): {last_error()}")
