"""ctypes binding of the C ABI in include/serenade_b200.h (libserenade_b200.so, built in-tree).

There is no fallback: if the shared library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_PKG = os.path.dirname(os.path.abspath(__file__))
# SRN_LIB_VARIANT=<name> loads an experiment build (python -m serenade_b200.build --variant <name> -DX=1)
_VARIANT = os.environ.get("SRN_LIB_VARIANT", "")
LIB_PATH = os.path.join(_PKG, f"libserenade_b200_{_VARIANT}.so" if _VARIANT else "libserenade_b200.so")

EXPORTED_SYMBOLS = [
    "srn_last_error", "srn_version", "srn_create", "srn_destroy", "srn_load_weight", "srn_finalize_weights",
    "srn_estimator_forward", "srn_solve_euler", "srn_solve_euler_host", "srn_launch_count", "srn_set_use_graph",
    "srn_debug_enable", "srn_debug_fetch", "srn_profile_ops", "srn_content_encode", "srn_style_encode",
    "srn_set_gn_valid_only", "srn_vocoder_set_dilations", "srn_vocoder_hop", "srn_vocoder_decode",
    "srn_estimator_forward_t", "srn_estimator_forward_tdev", "srn_solve_euler_tdev", "srn_cfm_loss_tdev", "srn_cfm_loss", "srn_set_option", "srn_softmax_stats",
    "srn_cfm_loss_from_velocity_noalloc",
]

PRECISION_BF16 = 0
PRECISION_TF32 = 1


class SrnConfig(C.Structure):
    _fields_ = [("in_channels", C.c_int32), ("out_channels", C.c_int32), ("spk_embed_dim", C.c_int32),
                ("channels", C.c_int32), ("num_heads", C.c_int32), ("attention_head_dim", C.c_int32),
                ("num_mid_blocks", C.c_int32), ("precision", C.c_int32)]


_lib: Optional[C.CDLL] = None


def load_library() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m serenade_b200.build` (needs nvcc; sm_100a only). "
            "serenade_b200 has no CPU or PyTorch fallback path.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, f32p, u8p = C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p
    lib.srn_last_error.restype = C.c_char_p
    lib.srn_last_error.argtypes = []
    lib.srn_version.restype = C.c_int
    lib.srn_create.argtypes = [C.POINTER(SrnConfig), C.POINTER(vp)]
    lib.srn_destroy.argtypes = [vp]
    lib.srn_load_weight.argtypes = [vp, C.c_char_p, f32p, i32, C.POINTER(i64)]
    lib.srn_finalize_weights.argtypes = [vp]
    lib.srn_estimator_forward.argtypes = [vp, f32p, f32p, u8p, f32p, C.c_float, i32, i32, f32p, vp]
    lib.srn_estimator_forward_t.argtypes = [vp, f32p, f32p, u8p, f32p, C.POINTER(C.c_float), i32, i32, f32p, vp]
    if hasattr(lib, "srn_estimator_forward_tdev"):
        lib.srn_estimator_forward_tdev.argtypes = [vp, f32p, f32p, u8p, f32p, f32p, i32, i32, i32, f32p, vp]
    lib.srn_cfm_loss.argtypes = [vp, f32p, f32p, C.POINTER(C.c_float), f32p, u8p, f32p, u8p, i32, i32, C.c_float, f32p, f32p,
                                 vp]
    if hasattr(lib, "srn_cfm_loss_tdev"):
        lib.srn_cfm_loss_tdev.argtypes = [vp, f32p, f32p, f32p, f32p, u8p, f32p, u8p, i32, i32, C.c_float, f32p, f32p, vp]
    lib.srn_solve_euler.argtypes = [vp, f32p, f32p, u8p, f32p, C.POINTER(C.c_float), i32, i32, i32, u8p, f32p, vp]
    if hasattr(lib, "srn_solve_euler_tdev"):
        lib.srn_solve_euler_tdev.argtypes = [vp, f32p, f32p, u8p, f32p, f32p, i32, i32, i32, u8p, f32p, vp]
    lib.srn_solve_euler_host.argtypes = [vp, f32p, f32p, u8p, f32p, C.POINTER(C.c_float), i32, i32, i32, f32p, vp]
    if hasattr(lib, "srn_cfm_loss_from_velocity_noalloc"):
        lib.srn_cfm_loss_from_velocity_noalloc.argtypes = [f32p, f32p, f32p, u8p, u8p, i32, i32, i32, C.c_float, f32p, vp]
    lib.srn_launch_count.restype = i64
    lib.srn_launch_count.argtypes = [vp]
    lib.srn_set_use_graph.argtypes = [vp, i32]
    lib.srn_set_gn_valid_only.argtypes = [vp, i32]
    if hasattr(lib, "srn_set_option"):      # (absent only from older experiment builds loaded with SRN_LIB_VARIANT)
        lib.srn_set_option.argtypes = [vp, C.c_char_p, i32]
        lib.srn_softmax_stats.argtypes = [vp, C.POINTER(i64)]
    lib.srn_vocoder_set_dilations.argtypes = [vp, C.POINTER(C.c_int32), i32]
    lib.srn_vocoder_hop.argtypes = [vp, C.POINTER(C.c_int32)]
    lib.srn_vocoder_decode.argtypes = [vp, f32p, f32p, f32p, i32, i32, f32p, vp]
    lib.srn_debug_enable.argtypes = [vp, i32]
    lib.srn_debug_fetch.argtypes = [vp, C.c_char_p, f32p, i64, C.POINTER(i64), C.POINTER(i32), C.POINTER(i32)]
    lib.srn_profile_ops.argtypes = [vp, i32, i32, i32, C.c_char_p, i64]
    lib.srn_content_encode.argtypes = [vp, f32p, i32, i32, f32p, vp]
    lib.srn_style_encode.argtypes = [vp, f32p, i32, i32, f32p, vp]
    for name in EXPORTED_SYMBOLS:
        if _VARIANT and not hasattr(lib, name):
            continue
        fn = getattr(lib, name)
        if name not in ("srn_last_error", "srn_launch_count", "srn_version"):
            fn.restype = C.c_int
    _lib = lib
    return lib


def check(status: int, what: str) -> None:
    if status != 0:
        msg = load_library().srn_last_error()
        raise RuntimeError(f"{what} failed: {msg.decode(errors='replace') if msg else 'unknown error'}")
