"""ctypes binding of libb200gru.so (C ABI in include/b200gru.h)."""
from __future__ import annotations

import ctypes
import os
import threading
from ctypes import (POINTER, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t,
                    c_uint32, c_uint64, c_void_p)

from . import _build
from .config import CConfig

_lock = threading.Lock()
_lib = None


class NativeError(RuntimeError):
    pass


class CIo(ctypes.Structure):
    """Mirror of `b200gru_io`."""

    _fields_ = [("rfloats", c_void_p), ("output", c_void_p),
                ("done_flag", c_void_p), ("done_value", c_uint32),
                ("ready_flag", c_void_p), ("ready_value", c_uint32),
                ("ready_event", c_void_p)]


def _declare(lib):
    P = POINTER
    cfg = P(CConfig)
    sig = {
        "b200gru_last_error": (c_char_p, []),
        "b200gru_version": (c_char_p, []),
        "b200gru_num_tensors": (c_int, [cfg]),
        "b200gru_param_count": (c_int64, [cfg]),
        "b200gru_param_layout": (c_int, [cfg, P(c_int64), P(c_int64), P(c_int32), P(c_int32)]),
        "b200gru_tensor_name": (c_char_p, [cfg, c_int]),
        "b200gru_read_binary": (c_int, [c_char_p, P(c_void_p), P(c_size_t)]),
        "b200gru_write_binary": (c_int, [c_char_p, c_void_p, c_size_t]),
        "b200gru_free_host": (None, [c_void_p]),
        "b200gru_random_init": (c_int, [cfg, c_uint64, c_void_p, c_int64]),
        "b200gru_random_floats": (c_int, [c_uint64, c_void_p, c_int64]),
        "b200gru_partition": (None, [c_int, c_int, c_int, P(c_int), P(c_int), P(c_int)]),
        "b200gru_random_select": (c_int, [c_void_p, c_int, c_float]),
        "b200gru_oracle_generate": (c_int, [cfg, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_uint32]),
        "b200gru_oracle_matvec": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int]),
        "b200gru_oracle_softmax": (c_int, [c_void_p, c_void_p, c_int, c_uint32]),
        "b200gru_oracle_gru_cell": (c_int, [c_void_p] * 6 + [c_int, c_int, c_uint32]),
        "b200gru_model_create": (c_int, [cfg, c_void_p, c_int, c_int, P(c_void_p)]),
        "b200gru_model_create_tp": (c_int, [cfg, c_void_p, c_int, c_int, c_int, c_int, P(c_void_p)]),
        "b200gru_tp_shared_bytes": (c_size_t, [c_void_p]),
        "b200gru_tp_bind": (c_int, [c_void_p, c_void_p, P(c_void_p)]),
        "b200gru_tp_bind_mc": (c_int, [c_void_p, c_void_p, P(c_void_p), c_void_p]),
        "b200gru_model_destroy": (c_int, [c_void_p]),
        "b200gru_model_device_bytes": (c_int64, [c_void_p]),
        "b200gru_generate": (c_int, [c_void_p, c_int, c_int, P(CIo), c_void_p]),
        "b200gru_last_launch_count": (c_int, [c_void_p]),
        "b200gru_model_set_debug": (c_int, [c_void_p, c_int]),
        "b200gru_model_set_path": (c_int, [c_void_p, c_int, c_int]),
        "b200gru_model_set_compaction": (c_int, [c_void_p, c_int]),
        "b200gru_debug_build": (c_int, []),
        "b200gru_model_check": (c_int, [c_void_p]),
        "b200gru_model_persist_timeline": (c_int, [c_void_p, c_void_p, c_int]),
        "b200gru_model_set_trace": (c_int, [c_void_p, c_int]),
        "b200gru_model_get_trace": (c_int, [c_void_p, P(c_float), P(c_char_p), c_int]),
        "b200gru_debug_logits": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
        "b200gru_wait_flags": (c_int, [c_void_p, c_int, c_uint32, c_void_p]),
        "b200gru_symm_alloc": (c_int, [c_int, c_size_t, P(c_void_p), c_void_p]),
        "b200gru_symm_open": (c_int, [c_int, c_void_p, P(c_void_p)]),
        "b200gru_symm_close": (c_int, [c_int, c_void_p]),
        "b200gru_symm_free": (c_int, [c_int, c_void_p]),
        "b200gru_memcpy_async": (c_int, [c_void_p, c_void_p, c_size_t, c_int, c_void_p]),
        "b200gru_memset_async": (c_int, [c_void_p, c_int, c_size_t, c_void_p]),
        "b200gru_signal_flag": (c_int, [c_void_p, c_uint32, c_void_p]),
        "b200gru_can_access_peer": (c_int, [c_int, c_int]),
        "b200gru_stream_synchronize": (c_int, [c_void_p]),
        "b200gru_device_count": (c_int, []),
        "b200gru_gru_cell_fp32": (c_int, [c_void_p] * 7 + [c_int, c_int, c_int, c_void_p]),
        "b200gru_p2p_write_bench": (c_int, [P(c_void_p), c_int, c_void_p, ctypes.c_longlong, c_int,
                                            ctypes.c_longlong, c_int, c_void_p]),
        "b200gru_gemm_bf16_tn": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    }
    for name, (res, args) in sig.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:  # symbol not built yet: fail at call time, loudly
            continue
        fn.restype = res
        fn.argtypes = args


def lib():
    """Load (building first if sources are newer) the native library."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if os.environ.get("B200GRU_NO_BUILD") != "1" and _build.needs_build():
                _build.build()
            if not _build.LIB_PATH.exists():
                raise NativeError(f"{_build.LIB_PATH} is missing; run `python -m gru_mpi_cuda_b200._build`")
            handle = ctypes.CDLL(str(_build.LIB_PATH), mode=ctypes.RTLD_GLOBAL)
            _declare(handle)
            _lib = handle
    return _lib


def check(rc: int, what: str = "") -> int:
    if rc < 0:
        msg = lib().b200gru_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'native call'} failed ({rc}): {msg}")
    return rc
