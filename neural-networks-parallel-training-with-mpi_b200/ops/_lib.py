"""ctypes binding of ``libb200dp.so`` (C ABI in ``include/b200dp.h``).

The library is built in-tree by ``build.py`` (nvcc, sm_100a).  ``lib()`` raises if it is missing:
on a GPU box there is no silent fallback to PyTorch ops.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Optional

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("B200DP_LIB") or os.path.join(_PKG, "lib", "libb200dp.so")   # override: A/B builds
_lock = threading.Lock()
_lib: Optional[C.CDLL] = None

c_void_p, c_int, c_ll, c_float = C.c_void_p, C.c_int, C.c_longlong, C.c_float
MAX_RANKS = 8
MAX_LAYERS = 64


class GemmArgs(C.Structure):
    _fields_ = [
        ("M", c_int), ("N", c_int), ("K", c_int),
        ("A", c_void_p), ("lda", c_ll), ("a_mn_major", c_int),
        ("B", c_void_p), ("ldb", c_ll), ("b_mn_major", c_int),
        ("D", c_void_p), ("ldd", c_ll), ("d_fp32", c_int),
        ("bias", c_void_p), ("relu", c_int),
        ("mask_src", c_void_p), ("ldmask", c_ll),
        ("tile_n", c_int), ("max_ctas", c_int), ("cta_group", c_int), ("split_tail", c_int),
        ("colsum_out", c_void_p), ("colsum_out_bf16", c_int),
        ("colsum_scratch", c_void_p), ("colsum_scratch_bytes", c_ll),
    ]


class B200dpError(RuntimeError):
    pass


def _declare(lib: C.CDLL) -> None:
    lib.b200dp_last_error.restype = C.c_char_p
    lib.b200dp_version.restype = c_int
    lib.b200dp_gemm_bf16.argtypes = [C.POINTER(GemmArgs), c_void_p]
    lib.b200dp_gemm_tf32.argtypes = [C.POINTER(GemmArgs), c_void_p]
    lib.b200dp_gemm_colsum_scratch_bytes.argtypes = [c_int, c_int]
    lib.b200dp_gemm_colsum_scratch_bytes.restype = c_ll
    for name, argtypes in _SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.argtypes = argtypes
            fn.restype = c_int


# filled in by the op modules' signatures below (kept in one table so the ABI is visible at a glance)
_SIGNATURES = {}


def register(name: str, argtypes) -> None:
    _SIGNATURES[name] = argtypes
    if _lib is not None and hasattr(_lib, name):
        getattr(_lib, name).argtypes = argtypes
        getattr(_lib, name).restype = c_int


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise B200dpError(
                        f"{LIB_PATH} is missing - run `python build.py` (nvcc, sm_100a). "
                        "The CUDA backend does not fall back to PyTorch ops.")
                l = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
                _declare(l)
                _lib = l
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise B200dpError(lib().b200dp_last_error().decode(errors="replace"))


def stream_ptr(stream=None) -> int:
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def ptr(t) -> Optional[int]:
    return None if t is None else t.data_ptr()


# launch counter: every call that enqueues one of OUR kernels bumps it (bench.py reports it)
launches = 0


def count(n: int = 1) -> None:
    global launches
    launches += n
