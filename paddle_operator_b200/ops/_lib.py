"""ctypes binding of libpdjob_b200.so (C ABI in include/pdjob_b200.h).

The library is loaded lazily.  On a machine with a GPU a missing or unloadable
library is a hard error: the ops in this package never fall back to eager
PyTorch on CUDA tensors.
"""
from __future__ import annotations

import ctypes
import os
import threading
from ctypes import POINTER, c_char_p, c_float, c_int, c_size_t, c_void_p

MAX_RANKS = 8
PAD_BYTES = 65536
MAX_BLOCKS = 1024        # barrier slots in the signal pad (csrc/common.cuh)
CHANNEL_BLOCKS = 256     # slots per channel; concurrent comm kernels must use different channels
NUM_CHANNELS = MAX_BLOCKS // CHANNEL_BLOCKS

AR_TWOSHOT, AR_ONESHOT, AR_NVLS = 0, 1, 2


class PdjComm(ctypes.Structure):
    _fields_ = [
        ("nranks", c_int),
        ("rank", c_int),
        ("pad_base", c_int),
        ("spin_timeout_ms", c_int),
        ("base", c_void_p * MAX_RANKS),
        ("mc_base", c_void_p),
    ]


class PdjEmbTable(ctypes.Structure):
    _fields_ = [
        ("deep_off", c_size_t), ("wide_off", c_size_t), ("gdeep_off", c_size_t), ("gwide_off", c_size_t),
        ("acc_deep_off", c_size_t), ("acc_wide_off", c_size_t), ("flag_off", c_size_t), ("list_off", c_size_t),
        ("count_off", c_size_t), ("rows_per_shard", ctypes.c_longlong), ("list_cap", ctypes.c_uint),
        ("dim", c_int), ("has_wide", c_int), ("shadow_off", c_size_t),
    ]


class PdjError(RuntimeError):
    pass


_lock = threading.Lock()
_lib = None


def _declare(lib: ctypes.CDLL) -> None:
    vp, sz, st = c_void_p, c_size_t, c_void_p
    fp = POINTER(c_float)
    sig = {
        "pdj_last_error": (c_char_p, []),
        "pdj_device_count": (c_int, []),
        "pdj_can_access_peer": (c_int, [c_int, c_int]),
        "pdj_multicast_supported": (c_int, [c_int]),
        "pdj_arena_round_size": (c_int, [c_int, sz, c_int, POINTER(sz)]),
        "pdj_arena_create": (c_int, [c_int, sz, POINTER(vp)]),
        "pdj_arena_export_fd": (c_int, [vp, POINTER(c_int)]),
        "pdj_arena_import_fd": (c_int, [c_int, c_int, sz, POINTER(vp)]),
        "pdj_arena_ptr": (vp, [vp]),
        "pdj_arena_bytes": (sz, [vp]),
        "pdj_arena_destroy": (c_int, [vp]),
        "pdj_multicast_create": (c_int, [c_int, c_int, sz, POINTER(vp), POINTER(c_int)]),
        "pdj_multicast_import": (c_int, [c_int, c_int, sz, POINTER(vp)]),
        "pdj_multicast_add_device": (c_int, [vp]),
        "pdj_multicast_bind": (c_int, [vp, vp, POINTER(vp)]),
        "pdj_multicast_destroy": (c_int, [vp]),
        "pdj_barrier": (c_int, [POINTER(PdjComm), c_int, st]),
        "pdj_allreduce_bf16": (c_int, [POINTER(PdjComm), sz, sz, sz, c_float, c_int, c_int, c_int, c_int, st]),
        "pdj_allreduce_sgd": (c_int, [POINTER(PdjComm), sz, sz, sz, sz, sz, sz, sz, vp, c_float, c_int, c_int, c_int, st]),
        "pdj_alltoall": (c_int, [POINTER(PdjComm), sz, sz, sz, c_int, c_int, st]),
        "pdj_broadcast": (c_int, [POINTER(PdjComm), c_int, sz, sz, c_int, c_int, st]),
        "pdj_emb_lookup": (c_int, [POINTER(PdjComm), POINTER(PdjEmbTable), sz, sz, sz, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_emb_push": (c_int, [POINTER(PdjComm), POINTER(PdjEmbTable), sz, sz, sz, c_int, c_int, c_int, c_float, c_float,
                                 c_int, c_int, st]),
        "pdj_emb_lookup2": (c_int, [POINTER(PdjComm), POINTER(PdjEmbTable), sz, sz, sz, c_int, c_int, c_int, c_int, sz, c_int,
                                    c_int, c_int, st]),
        "pdj_emb_push2": (c_int, [POINTER(PdjComm), POINTER(PdjEmbTable), sz, sz, sz, c_int, c_int, c_int, c_float, c_float,
                                  c_int, c_int, c_int, st]),
        "pdj_emb_apply": (c_int, [POINTER(PdjComm), POINTER(PdjEmbTable), c_int, c_float, c_float, c_int, c_int, st]),
        "pdj_linear_block_n": (c_int, [c_int]),
        "pdj_linear_bf16": (c_int, [vp, vp, vp, vp, vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_linear_bf16_kn": (c_int, [vp, vp, vp, vp, vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_conv1x1_partial_rows": (c_int, [c_int, c_int]),
        "pdj_conv1x1_bf16": (c_int, [vp, vp, vp, vp, c_int, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_conv1x1_add_bf16": (c_int, [vp, vp, vp, vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_conv1x1_kn_bf16": (c_int, [vp, vp, vp, vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_stem_s2d_partial_rows": (c_int, [c_int, c_int]),
        "pdj_stem_s2d_fwd": (c_int, [vp, vp, vp, vp, c_int, c_int, c_int, st]),
        "pdj_stem_s2d_wgrad_ws_bytes": (c_int, []),
        "pdj_stem_s2d_wgrad": (c_int, [vp, vp, vp, vp, c_int, c_int, c_int, st]),
        "pdj_conv_im2col_partial_rows": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, c_int]),
        "pdj_conv_im2col_fwd": (c_int, [vp, vp, vp, vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_conv3x3_wgrad_ws_bytes": (sz, [c_int, c_int, c_int, c_int, c_int]),
        "pdj_conv3x3_wgrad": (c_int, [vp, vp, vp, vp, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_wgrad_mn_ws_bytes": (sz, [c_int, c_int, c_int]),
        "pdj_wgrad_mn": (c_int, [vp, vp, vp, vp, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_wgrad_mn_db_ws_bytes": (sz, [c_int, c_int, c_int]),
        "pdj_wgrad_mn_db": (c_int, [vp, vp, vp, vp, vp, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_bn_fwd_pre": (c_int, [vp, vp, vp, vp, c_int, c_int, vp, vp, c_float, c_float, vp, vp, vp, vp, c_int, vp, c_int, st]),
        "pdj_bn_workspace_bytes": (sz, []),
        "pdj_bn_fwd": (c_int, [vp, vp, vp, c_int, c_int, vp, vp, c_float, c_float, vp, vp, vp, vp, c_int, vp, st]),
        "pdj_bn_bwd": (c_int, [vp, vp, vp, vp, vp, c_int, c_int, vp, vp, vp, vp, vp, c_int, vp, st]),
        "pdj_maxpool_fwd": (c_int, [vp, vp, vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_maxpool_bwd": (c_int, [vp, vp, vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, st]),
        "pdj_bn_fwd_gate": (c_int, [vp, vp, vp, vp, c_int, c_int, vp, vp, c_float, c_float, vp, vp, vp, vp, c_int, vp, st]),
        "pdj_bn_bwd_gate": (c_int, [vp, vp, vp, vp, vp, vp, c_int, c_int, vp, vp, vp, vp, vp, c_int, vp, st]),
        "pdj_bn_bwd_gate2": (c_int, [vp, vp, vp, vp, vp, vp, c_int, c_int, vp, vp, vp, vp, vp, vp, c_int, vp, st]),
        "pdj_sgd_momentum": (c_int, [vp, vp, vp, vp, sz, sz, vp, st]),
        "pdj_preprocess_u8": (c_int, [vp, vp, sz, fp, fp, st]),
        "pdj_preprocess_u8_s2d": (c_int, [vp, vp, c_int, c_int, c_int, fp, fp, st]),
        "pdj_wd_head": (c_int, [vp, c_int, vp, vp, vp, vp, vp, vp, vp, vp, vp, c_int, c_int, c_int, c_float, st]),
        "pdj_wd_head_dz": (c_int, [vp, vp, vp, c_int, vp, c_int, c_int, c_int, st]),
        "pdj_colsum_bf16": (c_int, [vp, vp, c_int, c_int, c_int, st]),
        "pdj_colsum_w_ws_bytes": (c_int, [c_int, c_int, c_int, c_int]),
        "pdj_colsum_w": (c_int, [vp, c_int, c_int, c_int, vp, vp, vp, c_int, c_int, c_int, vp, POINTER(c_void_p),
                                 POINTER(c_int), POINTER(c_int), POINTER(c_int), c_int, st]),
        "pdj_softmax_xent": (c_int, [vp, vp, vp, vp, c_int, c_int, c_float, st]),
    }
    for name, (res, args) in sig.items():
        if not hasattr(lib, name):
            continue  # optional entry points are declared by their own modules
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def lib_path() -> str:
    from ..build_ext import LIB_PATH

    return str(LIB_PATH)


def load(build_if_missing: bool = True) -> ctypes.CDLL:
    """Returns the loaded library, building it with nvcc first when needed."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        from .. import build_ext

        path = str(build_ext.LIB_PATH)
        if not os.path.exists(path) or (build_if_missing and not build_ext.is_fresh() and build_ext.shutil.which("nvcc")):
            if not build_if_missing:
                raise PdjError(f"{path} is missing; run `python -m paddle_operator_b200.build_ext`")
            build_ext.build()
        lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
        _declare(lib)
        _lib = lib
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().pdj_last_error()
        raise PdjError(f"{what or 'pdjob_b200 call'} failed (rc={rc}): {msg.decode() if msg else ''}")


def current_stream_ptr() -> int:
    import torch

    return torch.cuda.current_stream().cuda_stream


def ptr(t) -> int:
    """data_ptr of a tensor, or 0 for None."""
    return 0 if t is None else t.data_ptr()
