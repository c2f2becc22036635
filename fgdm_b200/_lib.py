"""ctypes binding of libfgdm_b200.so (C ABI declared in include/fgdm_b200.h).

The library is the product; this module only loads it and declares argument types. There is no
fallback: if the shared library is missing, or a compute entry point is called without an sm_100a
GPU, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FGDM_B200_LIB") or os.path.join(_HERE, "libfgdm_b200.so")   # override: A/B builds of one kernel

ABI_VERSION = 6

OUT_BF16, OUT_GEGLU, OUT_QKV, OUT_F32, OUT_NCHW_F32 = 0, 1, 2, 3, 4
ACT_NONE, ACT_SILU, ACT_RELU, ACT_QUICK_GELU = 0, 1, 2, 3
CONV_1X1, CONV_3X3, CONV_3X3_S2 = 0, 1, 2
CONV_UP2_P0 = 3          # + 2a + b: sub-pixel phase (a, b) of nearest-2x + conv3x3 (include/fgdm_b200.h)

vp, fp, ip = C.c_void_p, C.c_void_p, C.c_void_p  # all device pointers travel as integers


class GemmArgs(C.Structure):
    _fields_ = [
        ("a0", vp), ("a1", vp), ("c0", C.c_int), ("c1", C.c_int),
        ("nb", C.c_int), ("h", C.c_int), ("w", C.c_int), ("conv", C.c_int),
        ("weight", vp), ("n", C.c_int),
        ("bias", vp), ("rowvec", vp), ("ld_rowvec", C.c_int), ("rows_per_img", C.c_int),
        ("res0", vp), ("ld_res0", C.c_int), ("res1", vp), ("ld_res1", C.c_int),
        ("alpha", C.c_float), ("act", C.c_int), ("out_mode", C.c_int),
        ("out", vp), ("ldo", C.c_int), ("col_off", C.c_int),
        ("q_out", vp), ("k_out", vp), ("vt_out", vp), ("qkv_first", C.c_int),
        ("heads", C.c_int), ("dhead", C.c_int), ("tok_pad", C.c_int), ("bn_hint", C.c_int), ("n_store", C.c_int),
        ("res1_rows", C.c_int),
        ("row_stats", vp), ("ln_stats", vp), ("ln_slots", C.c_int), ("ln_s", vp), ("ln_channels", C.c_int),
        ("ln_eps", C.c_float),
    ]


# name -> argtypes (return type is always int except fgdm_last_error)
SIGNATURES = {
    "fgdm_abi_version": [],
    "fgdm_device_check": [],
    "fgdm_conv_gemm": [C.POINTER(GemmArgs), vp],
    "fgdm_gemm_stats_slots": [C.POINTER(GemmArgs)],
    "fgdm_attention": [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp],
    "fgdm_attention_causal": [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp],
    "fgdm_embed_tokens": [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp],
    "fgdm_groupnorm": [vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, vp],
    "fgdm_set_scratch_slot": [C.c_int],
    "fgdm_groupnorm_noalloc": [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, vp],
    "fgdm_layernorm": [vp, vp, vp, vp, C.c_int, C.c_int, C.c_float, vp],
    "fgdm_softmax_rows": [vp, vp, C.c_int, C.c_int, C.c_float, vp],
    "fgdm_cfg_ddim_step": [vp, vp, vp, vp, vp, vp, C.c_int, C.c_float, C.c_float, vp, vp, C.c_int, vp],
    "fgdm_plms_step": [vp, vp, vp, vp, vp, C.POINTER(C.c_float), vp, vp, C.c_int, C.c_float, vp, vp, vp, C.c_int, vp],
    "fgdm_timestep_embedding": [vp, vp, vp, C.c_int, C.c_int, vp],
    "fgdm_timestep_embedding_f32": [vp, vp, vp, C.c_int, C.c_int, vp],
    "fgdm_lincomb3": [C.c_float, vp, C.c_float, vp, C.c_float, vp, vp, C.c_int, vp],
    "fgdm_skinny_linear": [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp],
    "fgdm_resample": [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp],
    "fgdm_add_bf16": [vp, vp, vp, C.c_size_t, vp],
    "fgdm_conv_small_cin": [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int,
                            vp],
    "fgdm_im2col_small": [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, vp],
    "fgdm_ddim_advance": [vp, vp, vp, vp, C.c_int, vp, vp],
    "fgdm_conv_direct": [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp],
    "fgdm_conv_small_cout": [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp],
    "fgdm_latent_affine": [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_float, vp],
    "fgdm_image_to_u8": [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp],
    "fgdm_resize_u8": [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp],
    "fgdm_u8_to_hint": [vp, vp, C.c_int, C.c_int, C.c_int, vp],
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library (building nothing: see fgdm_b200.build). Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: run `python -m fgdm_b200.build` (nvcc, sm_100a). "
            "fgdm_b200 has no PyTorch/CPU fallback path.")
    lib = C.CDLL(LIB_PATH)
    lib.fgdm_last_error.restype = C.c_char_p
    lib.fgdm_last_error.argtypes = []
    lib.fgdm_launch_count.restype = C.c_longlong
    lib.fgdm_launch_count.argtypes = []
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = C.c_int
        fn.argtypes = argtypes
    if lib.fgdm_abi_version() != ABI_VERSION:
        raise RuntimeError("libfgdm_b200.so ABI version mismatch")
    _lib = lib
    return lib


def last_error() -> str:
    return load().fgdm_last_error().decode()


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {last_error()}")


def require_device() -> None:
    """Fail loudly unless an sm_100-class GPU and the TMA driver entry point are available."""
    check(load().fgdm_device_check(), "fgdm_device_check")
