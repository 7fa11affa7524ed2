"""Loader for the host emulation harness (tests/host_emul/tb_emul.cpp) -- test infrastructure.

It runs the kernel bodies with 32 fibers of one host thread standing in for the lanes of a warp, so the kernel
logic can be compared with the oracle without a GPU.  Never imported by the product package."""
import ctypes as C
import os
import subprocess

import numpy as np

from textblaster_b200 import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "host_emul", "tb_emul.cpp")
# TB_EMUL_SANITIZE=1: an AddressSanitizer + UndefinedBehaviorSanitizer build of the same source (tests/test_emul_sanitized.py
# runs it in a child process with the sanitizer runtimes preloaded)
_SANITIZE = os.environ.get("TB_EMUL_SANITIZE") == "1"
_SO = os.path.join(_HERE, "host_emul", "libtb_emul_asan.so" if _SANITIZE else "libtb_emul.so")
_CSRC = os.path.join(_HERE, "..", "textblaster_b200", "csrc")
_lib = None


def lib():
    global _lib
    if _lib is None:
        deps = [_SRC] + [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cuh", ".h", ".inc"))]
        if not os.path.exists(_SO) or any(os.path.getmtime(d) > os.path.getmtime(_SO) for d in deps):
            flags = ["-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-fno-omit-frame-pointer"] if _SANITIZE else ["-O2"]
            subprocess.run(["g++", "-std=c++20"] + flags + ["-fPIC", "-shared", "-pthread", "-x", "c++", _SRC, "-o", _SO], check=True)
        L = C.CDLL(_SO)
        L.tbe_run_chain.restype = C.c_int
        L.tbe_run_chain.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.tbe_chunk_targets.restype = C.c_int64
        L.tbe_chunk_targets.argtypes = [C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_int64]
        _lib = L
    return _lib


def rep_stats(reset=True):
    """(word ends seen, candidates kept, documents) of the signature-pruned duplicate n-gram levels since the last reset"""
    w, i, p = C.c_uint64(), C.c_uint64(), C.c_uint64()
    lib().tbe_rep_stats(C.byref(w), C.byref(i), C.byref(p), 1 if reset else 0)
    return w.value, i.value, p.value


def rep_task_stats(reset=True):
    """(levels worked off by one warp of a CTA tier in registers, ... with a table of its own) since the last reset"""
    a, b = C.c_uint64(), C.c_uint64()
    lib().tbe_rep_task_stats(C.byref(a), C.byref(b), 1 if reset else 0)
    return a.value, b.value


def chunk_targets(total_bytes, chunk_bytes, device_input=False):
    """tb_chunk_targets (csrc/tb_hostutil.h): the sizes the host executor cuts a batch into"""
    out = np.zeros(4096, np.int64)
    n = lib().tbe_chunk_targets(total_bytes, chunk_bytes, 1 if device_input else 0, out.ctypes.data, len(out))
    return [int(x) for x in out[:n]]


class EmulResult:
    pass


def run_chain(steps, texts, decode_entities=False):
    """steps: config param objects; texts: list[str].  Returns keep, failed_step, stats[], contents."""
    keepalive = []
    arr = (_abi.tb_step_cfg * max(len(steps), 1))()
    types = []
    for i, s in enumerate(steps):
        arr[i] = s.to_c(keepalive)
        types.append(arr[i].type)
    enc = [t.encode("utf-8") for t in texts]
    blob = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8).copy()
    off = np.zeros(len(enc) + 1, np.int64)
    if enc:
        off[1:] = np.cumsum([len(e) for e in enc])
    n = len(enc)
    keep = np.zeros(n + 1, np.uint8)
    failed = np.zeros(n + 1, np.int8)
    stats = [np.zeros(n + 1, _abi.stats_dtype(t)) for t in types]
    sp = (C.c_void_p * max(len(steps), 1))(*[s.ctypes.data for s in stats])
    content = np.zeros(3 * len(blob) + 8 * n + 512, np.uint8)  # sentence-wise lines can outgrow the input
    coff = np.zeros(n + 2, np.int64)
    rc = lib().tbe_run_chain(C.addressof(arr), len(steps), blob.ctypes.data, off.ctypes.data, n, keep.ctypes.data,
                             failed.ctypes.data, C.addressof(sp), content.ctypes.data, coff.ctypes.data, 1 if decode_entities else 0)
    if rc not in (0, 100):
        raise RuntimeError(f"emulation failed rc={rc}")
    r = EmulResult()
    r.keep, r.failed_step, r.stats = keep[:n], failed[:n], [s[:n] for s in stats]
    r.contents = None
    if rc == 100:
        r.contents = [bytes(content[coff[i]:coff[i + 1]]).decode("utf-8") for i in range(n)]
    return r
