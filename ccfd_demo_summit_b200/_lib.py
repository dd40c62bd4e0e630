"""ctypes binding of libccfd_rf.so (declared in include/ccfd_rf.h).

There is no CPU fallback: if the shared object is missing or does not load,
importing a scoring entry point raises, loudly.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_uint8, c_uint32, c_void_p

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libccfd_rf.so")

OK, EINVAL, ECUDA, ENOMEM, EUNSUPPORTED, ESTATE = 0, -1, -2, -3, -4, -5
DEVICE_PTRS, ASYNC = 0x1, 0x2
NOALLOC_MAX_TREE_ROWS = 1 << 20  # CCFD_NOALLOC_MAX_TREE_ROWS in include/ccfd_rf.h
KERNEL_AUTO, KERNEL_STREAM, KERNEL_SPLIT, KERNEL_GLOBAL = 0x000, 0x100, 0x200, 0x300
KERNELS = {"auto": KERNEL_AUTO, "stream": KERNEL_STREAM, "split": KERNEL_SPLIT, "global": KERNEL_GLOBAL}


class ForestDesc(ctypes.Structure):
    _fields_ = [
        ("n_trees", c_int32), ("n_features", c_int32), ("n_classes", c_int32), ("reserved", c_int32),
        ("tree_offsets", c_void_p), ("feature", c_void_p), ("threshold", c_void_p), ("left", c_void_p),
        ("right", c_void_p), ("missing_go_to_left", c_void_p), ("value", c_void_p),
    ]


class ForestInfo(ctypes.Structure):
    _fields_ = [
        ("n_trees", c_int32), ("n_features", c_int32), ("n_classes", c_int32), ("max_depth", c_int32),
        ("n_nodes", c_int64), ("n_leaves", c_int64), ("model_bytes", c_int64), ("max_tree_bytes", c_int64),
        ("device", c_int32), ("stream_kernel_ok", c_int32),
    ]


# every symbol include/ccfd_rf.h declares: (name, restype, argtypes)
SYMBOLS = [
    ("ccfd_device_count", c_int, []),
    ("ccfd_last_error", c_char_p, []),
    ("ccfd_forest_create", c_int, [POINTER(ForestDesc), c_int, POINTER(c_void_p)]),
    ("ccfd_forest_destroy", c_int, [c_void_p]),
    ("ccfd_forest_pack", c_int, [POINTER(ForestDesc), c_void_p, c_int64, c_void_p, c_void_p, POINTER(c_int64)]),
    ("ccfd_forest_get_info", c_int, [c_void_p, POINTER(ForestInfo)]),
    ("ccfd_predict_proba", c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_uint32, c_void_p]),
    ("ccfd_predict_decision", c_int,
     [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_double, c_int, c_uint32, c_void_p]),
    ("ccfd_predict_proba_rows", c_int, [c_void_p, c_void_p, c_int32, c_int64, c_int64, c_void_p, c_uint32, POINTER(c_int32)]),
    ("ccfd_kernel_for", c_int, [c_void_p, c_int64, POINTER(c_uint32), POINTER(c_int32)]),
    ("ccfd_predict_proba_noalloc", c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    ("ccfd_session_create", c_int, [c_void_p, c_int64, c_int32, POINTER(c_void_p)]),
    ("ccfd_session_destroy", c_int, [c_void_p]),
    ("ccfd_session_buffers", c_int, [c_void_p, c_int32, POINTER(POINTER(c_float)), POINTER(POINTER(c_double))]),
    ("ccfd_session_submit", c_int, [c_void_p, c_int32, c_int64]),
    ("ccfd_session_wait", c_int, [c_void_p, c_int32]),
    ("ccfd_session_run", c_int, [c_void_p, c_int32, c_int64]),
    ("ccfd_session_score", c_int, [c_void_p, c_int32, c_void_p, c_int32, c_int64, c_int64, c_void_p, POINTER(c_int32)]),
    ("ccfd_batcher_create", c_int, [c_void_p, c_int64, POINTER(c_void_p)]),
    ("ccfd_batcher_destroy", c_int, [c_void_p]),
    ("ccfd_batcher_push", c_int, [c_void_p, c_void_p, c_int64, c_int64]),
    ("ccfd_batcher_flush", c_int, [c_void_p]),
    ("ccfd_batcher_poll", c_int, [c_void_p, c_int32, POINTER(c_int64), POINTER(c_int64), c_void_p, c_int64, POINTER(c_double)]),
    ("ccfd_batcher_replay", c_int, [c_void_p, c_void_p, c_int64, c_int64, c_double, c_int32]),
    ("ccfd_batcher_stats", c_int, [c_void_p, POINTER(c_int64), POINTER(c_int64)]),
    ("ccfd_cast_rows", c_int, [c_void_p, c_int32, c_int64, c_int64, c_int64, c_void_p, POINTER(c_int32)]),
    ("ccfd_format_proba", c_int, [c_void_p, c_int64, c_void_p, c_int64, POINTER(c_int64)]),
    ("ccfd_parse_rows", c_int, [c_char_p, c_int64, c_int32, c_int32, c_void_p, c_int64, POINTER(c_int64)]),
]

_lib = None


class CcfdError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libccfd_rf error {code}: {msg}")
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            # a fresh checkout (the .so is not tracked): compile it in-tree if nvcc is here; there is no
            # other way to score, so anything else is an error
            try:
                from . import build as _build

                _build.build()
            except Exception as e:  # noqa: BLE001
                raise ImportError(
                    f"{LIB_PATH} is missing and could not be built ({e}): run `python -m ccfd_demo_summit_b200.build` "
                    "(nvcc, sm_100a). There is no CPU fallback for the scoring path."
                ) from e
        handle = ctypes.CDLL(LIB_PATH)
        for name, restype, argtypes in SYMBOLS:
            fn = getattr(handle, name)  # AttributeError if the build lost a symbol
            fn.restype, fn.argtypes = restype, argtypes
        _lib = handle
    return _lib


def check(rc: int):
    """0 -> ok; CCFD_EINVAL -> ValueError (as sklearn raises for bad input); anything else -> RuntimeError."""
    if rc >= 0:
        return rc
    msg = (lib().ccfd_last_error() or b"").decode("utf-8", "replace")
    if rc == EINVAL:
        raise ValueError(msg)
    raise CcfdError(rc, msg)
