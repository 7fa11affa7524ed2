"""Synthetic C4-shaped corpus as an Arrow-style string column: uint8 values + int64
offsets.  Thin ctypes wrapper over csrc/tb_synth.cpp (host-only g++ helper, built in-tree)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "csrc", "tb_synth.cpp")
_SO = os.path.join(_HERE, "libtb_synth.so")
_lib = None

BASE_SEED = 0x7EB1A57E  # seed = 0x7EB1A57E + config_id


def build(force: bool = False) -> str:
    if force or not os.path.exists(_SO) or os.path.getmtime(_SRC) > os.path.getmtime(_SO):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", _SRC, "-o", _SO], check=True)
    return _SO


def _load():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.tbs_generate.restype = C.c_void_p
        L.tbs_generate.argtypes = [C.c_int64, C.c_uint64, C.c_int64, C.c_double, C.c_double, C.c_int, C.POINTER(C.c_int64)]
        L.tbs_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.tbs_free.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def generate(n_docs: int, seed: int = BASE_SEED, median_bytes: float = 2048.0, sigma: float = 1.0, first_index: int = 0,
             nthreads: int = 0, out=None):
    """Returns (utf8 uint8[total], offsets int64[n_docs + 1]).  ``out`` may be a callable
    ``(nbytes) -> uint8 ndarray`` used to allocate the values buffer (e.g. pinned memory)."""
    L = _load()
    total = C.c_int64()
    h = L.tbs_generate(n_docs, seed, first_index, median_bytes, sigma, nthreads or (os.cpu_count() or 1), C.byref(total))
    try:
        buf = out(total.value) if out is not None else np.empty(max(total.value, 1), np.uint8)
        off = np.empty(n_docs + 1, np.int64)
        L.tbs_copy(h, buf.ctypes.data, off.ctypes.data)
    finally:
        L.tbs_free(h)
    return buf[:total.value], off
