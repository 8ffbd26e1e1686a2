"""ctypes binding of libb200ps.so — the hand-written sm_100a kernels behind a C ABI.

The library is built in-tree by ``build.py`` (``b200ps/_lib/libb200ps.so``) so that it
travels with the repository snapshot to the GPU box.  On a machine with a CUDA device a
missing library is a hard error: the framework never silently falls back to PyTorch ops.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

# B200PS_CHECKED=1 selects the build with device-side protocol assertions (build.py --checked)
_CHECKED = os.environ.get("B200PS_CHECKED", "0") == "1"
_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_lib",
                         "libb200ps_checked.so" if _CHECKED else "libb200ps.so")
if _CHECKED:
    os.environ["B200PS_FASTFEED"] = "0"      # the CPython fast path is linked against the release library
_lib = None
_lock = threading.Lock()


class NativeError(RuntimeError):
    pass


class GemmArgs(C.Structure):
    """Mirror of ``b200ps_gemm_args`` (include/b200ps.h)."""
    _fields_ = [
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("ab_dtype", C.c_int32), ("a_mn_major", C.c_int32), ("b_mn_major", C.c_int32),
        ("A", C.c_void_p), ("lda", C.c_int64),
        ("B", C.c_void_p), ("ldb", C.c_int64),
        ("C", C.c_void_p), ("ldc", C.c_int64),
        ("c_dtype", C.c_int32), ("epilogue", C.c_int32),
        ("alpha", C.c_float),
        ("bias", C.c_void_p),
        ("act", C.c_int32),
        ("aux", C.c_void_p), ("ldaux", C.c_int64),
        ("split_k", C.c_int32),
        ("sgd_master", C.c_void_p), ("sgd_lowp", C.c_void_p),
        ("sgd_lr", C.c_float),
        ("max_ctas", C.c_int32),
        ("a_row_offset", C.c_void_p), ("a_rows_total", C.c_int64), ("c_split_stride", C.c_int64),
        ("gather_src", C.c_void_p), ("gather_flags", C.c_void_p),
        ("gather_world", C.c_int32), ("gather_rank", C.c_int32), ("gather_ctas", C.c_int32),
        ("colsum_out", C.c_void_p),
    ]


class FusedPush(C.Structure):
    """Mirror of ``b200ps_fused_push`` (include/b200ps.h)."""
    _fields_ = [
        ("pads", C.c_void_p), ("masters", C.c_void_p), ("bounds", C.c_void_p), ("ll_grads", C.c_void_p),
        ("ll_params", C.c_void_p), ("global_steps", C.c_void_p), ("max_shard_elems", C.c_int64),
        ("off_w1", C.c_int64), ("off_b1", C.c_int64), ("off_w2", C.c_int64), ("off_b2", C.c_int64),
        ("world", C.c_int32), ("rank", C.c_int32),
    ]


DT_BF16, DT_F32 = 0, 1
ACT_NONE, ACT_SIGMOID = 0, 1
EPI_STORE, EPI_ATOMIC_ADD, EPI_SGD = 0, 1, 2
TAIL_OVERWRITE_STATS, TAIL_CLEAR_Z2 = 1, 2


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def _declare(lib):
    vp, i32, i64, u64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_float
    lib.b200ps_last_error.restype = C.c_char_p
    lib.b200ps_version.restype = C.c_char_p
    lib.b200ps_launch_count.restype = u64
    sigs = {
        "b200ps_gemm": [C.POINTER(GemmArgs), vp],
    }
    from . import _signatures  # remaining entry points, kept next to their wrappers
    sigs.update(_signatures.SIGNATURES)
    for name, argtypes in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = C.c_int
    for name, (restype, argtypes) in _signatures.SPECIAL.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = restype


def load():
    """Load (once) and return the ctypes handle of libb200ps.so."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(_LIB_PATH):
                raise NativeError(
                    f"{_LIB_PATH} is missing: run `python build.py` (nvcc, sm_100a) first")
            lib = C.CDLL(_LIB_PATH, mode=C.RTLD_GLOBAL)
            _declare(lib)
            _lib = lib
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().b200ps_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'b200ps'} failed ({rc}): {msg}")


def launch_count() -> int:
    return int(load().b200ps_launch_count())


def version() -> str:
    return load().b200ps_version().decode()


def stream_ptr(stream=None) -> int:
    """Raw cudaStream_t of ``stream`` (default: torch's current stream)."""
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return int(s.cuda_stream)


def ptr(t) -> int:
    return 0 if t is None else int(t.data_ptr())
