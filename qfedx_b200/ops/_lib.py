"""ctypes loader for the in-tree CUDA library (qfedx_b200/lib/libqfedx_b200.so).

On a machine with a GPU the library is mandatory: ``lib()`` raises if it is missing or stale
instead of letting callers fall back to PyTorch silently.  On a CPU-only machine the ops use
their torch reference path and never call ``lib()``.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

import torch

_LIB = None
LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libqfedx_b200.so"
LAUNCH_COUNT = 0  # kernels launched through this module (bench.py reports it)


class QfxError(RuntimeError):
    pass


def available() -> bool:
    return LIB_PATH.exists()


def lib():
    global _LIB
    if _LIB is None:
        if not LIB_PATH.exists():
            raise QfxError(f"{LIB_PATH} is missing: run `python -m qfedx_b200.build` "
                           "(the CUDA path never falls back to PyTorch on a GPU machine)")
        _LIB = ctypes.CDLL(str(LIB_PATH), mode=ctypes.RTLD_GLOBAL)
        _LIB.qfx_last_error.restype = ctypes.c_char_p
        _LIB.qfx_version.restype = ctypes.c_int
    return _LIB


def use_cuda(*tensors) -> bool:
    """True when the op should run on the hand-written kernels."""
    return any(t is not None and t.is_cuda for t in tensors)


def ptr(t, dtype=None):
    if t is None:
        return ctypes.c_void_p(0)
    if dtype is not None and t.dtype != dtype:
        raise QfxError(f"expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise QfxError("tensor must be contiguous")
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def call(name: str, *args, launches: int = 1):
    """Invoke an entry point; raise with the library's message on a non-zero return."""
    global LAUNCH_COUNT
    fn = getattr(lib(), name)
    fn.restype = ctypes.c_int
    rc = fn(*args)
    if rc != 0:
        raise QfxError(f"{name} failed (code {rc}): {lib().qfx_last_error().decode()}")
    LAUNCH_COUNT += launches
    return rc


c_int = ctypes.c_int
c_float = ctypes.c_float
c_u64 = ctypes.c_ulonglong
c_u32 = ctypes.c_uint
