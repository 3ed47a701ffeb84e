"""ctypes binding of libperseus_b200.so (the C ABI declared in include/perseus_b200.h).

There is no fallback: if the library is missing this raises, and every compute entry point
returns PB200_ERR_NO_DEVICE (surfaced as Pb200Error) when no sm_100 GPU is present.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int64, c_size_t, c_uint8, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libperseus_b200.so")

PREC_FP32, PREC_BF16, PREC_BF16X3, PREC_FP16X2 = 0, 1, 2, 3
PRECISIONS = {"fp32": PREC_FP32, "bf16": PREC_BF16, "bf16x3": PREC_BF16X3, "fp16x2": PREC_FP16X2}

ERR_NO_DEVICE = 3


class Pb200Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"perseus_b200 error {code}: {msg}")
        self.code = code


# every symbol include/perseus_b200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "pb200_version": (c_int, []),
    "pb200_last_error": (c_char_p, []),
    "pb200_plan_create": (c_int, [POINTER(c_void_p), c_int, c_int, c_int, c_int, c_int, c_int, c_int]),
    "pb200_plan_destroy": (None, [c_void_p]),
    "pb200_plan_workspace_bytes": (c_int, [c_void_p, POINTER(c_size_t)]),
    "pb200_plan_frames_bytes": (c_int, [c_void_p, c_int, POINTER(c_size_t)]),
    "pb200_plan_load_weights": (
        c_int,
        [c_void_p, POINTER(c_void_p), POINTER(POINTER(c_int64)), POINTER(c_int), POINTER(c_char_p), c_int, c_void_p],
    ),
    "pb200_ingest": (
        c_int,
        [c_void_p, c_void_p, c_int, c_int64, c_int, c_int, c_int, c_void_p, c_float, c_int, c_float, c_float,
         c_float, c_void_p, c_void_p, c_int, c_void_p],
    ),
    "pb200_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "pb200_forward_rgb8": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "pb200_forward_nchw_f32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "pb200_plan_num_ops": (c_int, [c_void_p]),
    "pb200_plan_op_name": (c_char_p, [c_void_p, c_int]),
    "pb200_plan_op_kernel": (c_char_p, [c_void_p, c_int]),
    "pb200_plan_op_work": (c_int, [c_void_p, c_int, c_int, POINTER(c_double), POINTER(c_double)]),
    "pb200_forward_profile": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, POINTER(c_float)]
    ),
    "pb200_plan_launches_per_forward": (c_int, [c_void_p]),
    "pb200_plan_launches_at": (c_int, [c_void_p, c_int]),
    "pb200_plan_op_chain_len": (c_int, [c_void_p, c_int]),
    "pb200_decode_pixels": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "pb200_debug_conv_bf16": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int,
         c_void_p, c_void_p],
    ),
    "pb200_debug_conv_ds_bf16": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p],
    ),
    "pb200_debug_conv_split": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int,
         c_void_p, c_void_p],
    ),
    "pb200_debug_stem_pool": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "pb200_debug_conv_p8": (
        c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p]
    ),
    "pb200_debug_conv_f16x2": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int,
         c_void_p, c_void_p, c_int, c_void_p, c_void_p],
    ),
    "pb200_debug_conv_halo": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p,
         c_void_p],
    ),
    "pb200_debug_conv_p8_f16x2": (
        c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p]
    ),
    "pb200_train_pack_weights": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "pb200_train_pack_all": (c_int, [c_void_p, c_int, c_int64, c_void_p]),
    "pb200_train_conv": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int,
         c_void_p, c_void_p, c_void_p],
    ),
    "pb200_train_dgrad_s2": (
        c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]
    ),
    "pb200_train_wgrad": (
        c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]
    ),
    "pb200_train_unpack_wgrad": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "pb200_train_bn_stats": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p]),
    "pb200_train_bn_apply": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float, c_void_p, c_void_p, c_void_p,
         c_int, c_void_p, c_int64, c_int, c_void_p],
    ),
    "pb200_train_bn_backward": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p,
         c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p],
    ),
    "pb200_smoother_normal_equations": (
        c_int,
        [c_void_p] * 9 + [POINTER(c_double), c_double, c_double, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p],
    ),
    "pb200_smoother_retract": (c_int, [c_void_p] * 7 + [c_int64, c_void_p]),
    "pb200_smoother_push": (c_int, [c_void_p] * 5 + [c_double, c_int, c_int, c_int, c_int, c_void_p]),
    "pb200_smoother_solve": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "pb200_augment": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, ctypes.c_uint64, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
         c_int, c_int, c_int, c_int, c_int, c_void_p],
    ),
    "pb200_plasma_side": (c_int, [c_int, c_int]),
    "pb200_plasma": (c_int, [c_void_p, ctypes.c_uint64, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "pb200_plasma_static": (c_int, [c_void_p, ctypes.c_uint64, c_void_p, c_int, c_int, c_int, c_void_p]),
    "pb200_train_stem_im2col": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "pb200_train_stem_pack": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "pb200_train_stem_unpack": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "pb200_train_maxpool": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "pb200_train_maxpool_backward": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "pb200_train_head_forward": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]
    ),
    "pb200_train_smooth_l1": (c_int, [c_void_p, c_void_p, c_int, c_float, c_void_p, c_void_p, c_void_p]),
    "pb200_train_head_backward": (
        c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]
    ),
    "pb200_train_adamw_step": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float, c_float, c_float, c_float, c_int, c_float,
         c_float, c_void_p, c_void_p],
    ),
    "pb200_factors_projection": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p],
    ),
    "pb200_factors_dynamics": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_double, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
         c_int64, c_void_p],
    ),
    "pb200_factors_constvel": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "pb200_factors_chain": (
        c_int,
        [c_void_p] * 7 + [c_double, c_int] + [c_void_p] * 10 + [c_int, c_int, c_int, c_void_p],
    ),
}

_lib = None


def lib() -> ctypes.CDLL:
    """Loads the native library (once).  Raises if it has not been built — no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -m perseus_b200.build` "
                "(perseus_b200 has no CPU / PyTorch fallback path)"
            )
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise Pb200Error(rc, lib().pb200_last_error().decode("utf-8", "replace"))


def current_stream_ptr() -> int:
    import torch

    return torch.cuda.current_stream().cuda_stream


def ptr(t) -> int | None:
    """Device/host pointer of a torch tensor (None passes NULL)."""
    return None if t is None else t.data_ptr()
