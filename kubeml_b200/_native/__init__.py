"""Loader and ctypes signatures for the native runtime (csrc/, include/kubeml_b200.h).

The shared object is built in-tree by `kubeml_b200/_native/build.py` (plain nvcc, sm_100a only)
and travels with the source snapshot.  On a GPU box a missing or stale library is a hard error:
CUDA ops never fall back silently.
"""
from __future__ import annotations

import ctypes
import os
import threading
from ctypes import (POINTER, Structure, c_char_p, c_float, c_int, c_longlong, c_size_t, c_uint32,
                    c_ulonglong, c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("KUBEML_NATIVE_LIB") or os.path.join(_HERE, "libkubeml_b200.so")     # override: A/B two builds
KML_MAX_RANKS = 8

_lib = None
_lock = threading.Lock()


class NativeError(RuntimeError):
    pass


class MergeParams(Structure):
    _fields_ = [
        ("replica", c_void_p * KML_MAX_RANKS),
        ("ireplica", c_void_p * KML_MAX_RANKS),
        ("flags", c_void_p * KML_MAX_RANKS),
        ("mcast", c_void_p),
        ("n_f32", c_size_t),
        ("n_i64", c_size_t),
        ("n_sgd", c_size_t),
        ("my_rank", c_int),
        ("participants", c_uint32),
        ("write_mask", c_uint32),
        ("generation", c_uint32),
        ("fuse_sgd", c_int),
        ("grad", c_void_p),
        ("momentum_buf", c_void_p),
        ("lr", c_float),
        ("weight_decay", c_float),
        ("momentum", c_float),
        ("w_bf16", c_void_p),
        ("timeout_cycles", c_ulonglong),
        ("status", c_void_p),
        ("state2", c_void_p),
        ("beta1", c_float),
        ("beta2", c_float),
        ("eps", c_float),
        ("step_dev", c_void_p),
        ("n_fuse", c_size_t),
    ]


MERGE_EVICTED = 0x40000000      # KML_MERGE_EVICTED / KML_MERGE_LATE of include/kubeml_b200.h
MERGE_LATE = 0x20000000


class BnFuse(Structure):
    """kml_bn_fuse (include/kubeml_b200.h): a BatchNorm finished inside a cluster-mode GEMM launch."""
    _fields_ = [("mode", c_int), ("scratch", c_void_p), ("gamma", c_void_p), ("beta", c_void_p),
                ("running_mean", c_void_p), ("running_var", c_void_p), ("num_batches_tracked", c_void_p),
                ("save_mean", c_void_p), ("save_invstd", c_void_p), ("res", c_void_p), ("raw_out", c_void_p),
                ("relu", c_int), ("eps", c_float), ("momentum", c_float), ("mask_y", c_void_p), ("bn_x", c_void_p),
                ("dgamma", c_void_p), ("dbeta", c_void_p)]


class Epilogue(Structure):
    _fields_ = [("bias", c_void_p), ("addend_bf16", c_void_p), ("relu", c_int), ("out_f32", c_int),
                ("splits", c_int), ("slabs", c_int), ("block_n", c_int), ("cluster", c_int), ("stats", c_void_p),
                ("mask_y", c_void_p), ("bn_x", c_void_p), ("bn_mean", c_void_p), ("bn_invstd", c_void_p),
                ("bn_fuse", POINTER(BnFuse))]


def _declare(lib) -> None:
    def sig(name, restype, *argtypes):
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = list(argtypes)

    sig("kml_last_error", c_char_p)
    sig("kml_version", c_int)
    sig("kml_launch_count", c_ulonglong)
    sig("kml_reset_launch_count", None)
    sig("kml_num_sms", c_int)
    # arena
    sig("kml_arena_create", c_int, c_int, c_size_t, c_int, POINTER(c_void_p))
    sig("kml_arena_destroy", c_int, c_void_p)
    sig("kml_arena_ptr", c_void_p, c_void_p)
    sig("kml_arena_size", c_size_t, c_void_p)
    sig("kml_arena_export_fd", c_int, c_void_p, POINTER(c_int))
    sig("kml_arena_import_fd", c_int, c_int, c_int, c_size_t, POINTER(c_void_p))
    sig("kml_arena_grant_access", c_int, c_void_p, c_int)
    sig("kml_mcast_supported", c_int, c_int, POINTER(c_int))
    sig("kml_mcast_create", c_int, c_int, c_size_t, POINTER(c_void_p))
    sig("kml_mcast_export_fd", c_int, c_void_p, POINTER(c_int))
    sig("kml_mcast_import_fd", c_int, c_int, c_int, c_size_t, POINTER(c_void_p))
    sig("kml_mcast_add_device", c_int, c_void_p, c_int)
    sig("kml_mcast_bind", c_int, c_void_p, c_int, c_void_p, POINTER(c_void_p))
    sig("kml_mcast_destroy", c_int, c_void_p)
    # merge / optim / loss
    sig("kml_kavg_merge", c_int, POINTER(MergeParams), c_void_p)
    sig("kml_gemm_max_coresident", c_int, c_int, c_int, c_int, c_int)
    sig("kml_merge_mark_dead", c_int, c_void_p, c_int, c_uint32, c_uint32, c_void_p)
    sig("kml_sgd_step", c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_float, c_float, c_float,
        c_int, c_void_p, c_void_p)
    sig("kml_adam_step", c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_float, c_float,
        c_float, c_float, c_float, c_int, c_void_p, c_void_p)
    sig("kml_adam_step_dev", c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_float, c_float,
        c_float, c_float, c_float, c_void_p, c_void_p, c_void_p)
    sig("kml_cast_f32_to_bf16", c_int, c_void_p, c_void_p, c_size_t, c_void_p)
    sig("kml_fill_f32", c_int, c_void_p, c_float, c_size_t, c_void_p)
    sig("kml_softmax_ce_scratch_bytes", c_size_t, c_int)
    sig("kml_softmax_ce", c_int, c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p,
        c_void_p, c_void_p, c_void_p)
    # tcgen05 GEMM family
    sig("kml_gemm_bf16", c_int, c_void_p, c_int, c_longlong, c_void_p, c_int, c_longlong, c_void_p,
        c_longlong, c_int, c_int, c_int, POINTER(Epilogue), c_void_p)
    sig("kml_conv2d_fprop", c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
        c_int, c_int, c_int, POINTER(Epilogue), c_void_p)
    sig("kml_conv2d_fprop_strided", c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
        c_int, c_int, c_int, c_int, POINTER(Epilogue), c_void_p)
    sig("kml_conv2d_wgrad_strided", c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
        c_int, c_int, c_int, c_int, c_int, c_int, c_void_p)
    sig("kml_conv2d_dgrad", c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
        c_int, c_int, c_int, POINTER(Epilogue), c_void_p)
    sig("kml_conv2d_wgrad", c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
        c_int, c_int, c_int, c_int, c_int, c_void_p)
    # NHWC layers
    sig("kml_bn_scratch_bytes", c_size_t, c_int)
    sig("kml_bn_stats_bytes", c_size_t, c_int)
    sig("kml_bn_apply_train", c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
        c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_float, c_void_p)
    sig("kml_bn_bwd_apply", c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
        c_void_p, c_int, c_int, c_void_p)
    sig("kml_bn_fwd_train", c_int, c_void_p, c_void_p, c_int, c_longlong, c_void_p, c_void_p, c_void_p,
        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
        c_float, c_float, c_void_p)
    sig("kml_bn_fwd_eval", c_int, c_void_p, c_void_p, c_int, c_longlong, c_void_p, c_void_p, c_void_p, c_void_p,
        c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p)
    sig("kml_bn_bwd", c_int, c_void_p, c_void_p, c_int, c_longlong, c_void_p, c_void_p, c_void_p, c_void_p,
        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p)
    sig("kml_maxpool_fwd", c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
        c_int, c_void_p)
    sig("kml_maxpool_bwd", c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
        c_int, c_void_p)
    sig("kml_avgpool_fwd", c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p)
    sig("kml_avgpool_bwd", c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p)
    sig("kml_im2col", c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
        c_int, c_void_p)
    sig("kml_col2im", c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
        c_int, c_int, c_void_p)
    sig("kml_colsum_bf16", c_int, c_void_p, c_void_p, c_int, c_int, c_longlong, c_void_p)
    sig("kml_stamp", c_int, c_void_p, c_int, c_void_p)
    sig("kml_post_f32", c_int, c_void_p, c_void_p, c_int, c_void_p)
    sig("kml_image_prep_rand", c_int, c_void_p, c_void_p, c_void_p, c_uint32, c_int, c_int, c_int, c_int, c_int,
        POINTER(c_float), POINTER(c_float), c_void_p, c_void_p, c_void_p)
    sig("kml_pack_rows_bf16", c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p)
    sig("kml_relu_bwd", c_int, c_void_p, c_void_p, c_void_p, c_float, c_size_t, c_void_p)
    sig("kml_relu_bwd_cols", c_int, c_void_p, c_void_p, c_int, c_size_t, c_void_p, c_void_p, c_float, c_void_p,
        c_int, c_int, c_void_p)
    sig("kml_bias_act", c_int, c_void_p, c_int, c_size_t, c_void_p, c_void_p, c_int, c_int, c_int, c_longlong,
        c_longlong, c_void_p)
    sig("kml_dropout_fwd", c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_uint32, c_float, c_size_t, c_void_p)
    sig("kml_dropout_bwd", c_int, c_void_p, c_void_p, c_void_p, c_float, c_size_t, c_void_p)
    sig("kml_bump_counter", c_int, c_void_p, c_void_p)
    sig("kml_replicate_fwd", c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p)
    sig("kml_replicate_bwd", c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p)
    sig("kml_image_prep", c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
        POINTER(c_float), POINTER(c_float), c_void_p)
    for name, s in _EXTRA_SIGS.items():
        if hasattr(lib, name):
            sig(name, *s)


# filled by later modules of this package (norm, pool, ...) before the first load
_EXTRA_SIGS: dict = {}


def available() -> bool:
    return os.path.isfile(LIB_PATH)


def lib():
    """The loaded library; raises NativeError with a build hint if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.isfile(LIB_PATH):
                raise NativeError(
                    f"{LIB_PATH} not found: build it with `python -m kubeml_b200._native.build` "
                    "(or __graft_entry__.build()); CUDA ops do not fall back to PyTorch")
            l = ctypes.CDLL(LIB_PATH)
            _declare(l)
            _lib = l
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().kml_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg}")


def stream_ptr(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return int(s.cuda_stream)


def launch_count() -> int:
    return int(lib().kml_launch_count()) if _lib is not None or available() else 0
