"""Python loader for the CPU oracle (oracle/tb_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Nothing under textblaster_b200/ imports this package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from textblaster_b200 import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libtb_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "tb_oracle.cpp")
    deps = [src, os.path.join(_HERE, "..", "include", "textblaster_b200.h"),
            os.path.join(_HERE, "..", "textblaster_b200", "csrc", "unicode_tables.inc")]
    stale = force or not os.path.exists(_SO) or any(os.path.getmtime(d) > os.path.getmtime(_SO) for d in deps)
    if stale:
        subprocess.run(["make", "-C", _HERE, "-s", "-B", "libtb_oracle.so"], check=True)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.tbo_pipeline_create.restype = C.c_void_p
        L.tbo_pipeline_create.argtypes = [C.c_void_p, C.c_size_t]
        L.tbo_pipeline_destroy.argtypes = [C.c_void_p]
        L.tbo_run_batch.restype = C.c_void_p
        L.tbo_run_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int]
        L.tbo_result_free.argtypes = [C.c_void_p]
        for f in ("tbo_result_keep", "tbo_result_failed_step"):
            getattr(L, f).restype = C.c_void_p
            getattr(L, f).argtypes = [C.c_void_p]
        L.tbo_result_stats.restype = C.c_void_p
        L.tbo_result_stats.argtypes = [C.c_void_p, C.c_int]
        for f in ("tbo_result_reason", "tbo_result_metadata"):
            getattr(L, f).restype = C.c_char_p
            getattr(L, f).argtypes = [C.c_void_p, C.c_int64]
        L.tbo_result_content.restype = C.c_void_p
        L.tbo_result_content.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
        L.tbo_result_content_total.restype = C.c_int64
        L.tbo_result_content_total.argtypes = [C.c_void_p]
        L.tbo_result_content_packed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        for f in ("tbo_split_words", "tbo_split_sentences", "tbo_word_boundaries", "tbo_sentence_boundaries"):
            getattr(L, f).restype = C.c_int64
            getattr(L, f).argtypes = [C.c_char_p, C.c_int64, C.c_void_p, C.c_int64]
        L.tbo_find_duplicates.argtypes = [C.c_char_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.tbo_find_top_duplicate.restype = C.c_int64
        L.tbo_find_top_duplicate.argtypes = [C.c_char_p, C.c_void_p, C.c_int64]
        L.tbo_find_all_duplicate.restype = C.c_int64
        L.tbo_find_all_duplicate.argtypes = [C.c_char_p, C.c_void_p, C.c_int64, C.c_int64]
        L.tbo_get_n_grams.restype = C.c_int64
        L.tbo_get_n_grams.argtypes = [C.c_char_p, C.c_void_p, C.c_int64, C.c_int64, C.c_char_p, C.c_int64]
        L.tbo_punctuation_contains.argtypes = [C.c_uint32]
        L.tbo_cinfo.restype = C.c_uint32
        L.tbo_cinfo.argtypes = [C.c_uint32]
        for f in ("tbo_to_lowercase", "tbo_remove_citations", "tbo_decode_html_entities"):
            getattr(L, f).restype = C.c_int64
            getattr(L, f).argtypes = [C.c_char_p, C.c_int64, C.c_char_p, C.c_int64]
        _lib = L
    return _lib


def _pairs(fn, text: str):
    b = text.encode("utf-8")
    cap = len(b) + 2
    se = (C.c_int64 * (2 * cap))()
    n = fn(b, len(b), se, cap)
    return [b[se[2 * i]:se[2 * i + 1]].decode("utf-8") for i in range(n)]


def split_into_words(text: str):
    return _pairs(lib().tbo_split_words, text)


def split_into_sentences(text: str):
    return _pairs(lib().tbo_split_sentences, text)


def _bounds(fn, text):
    b = text.encode("utf-8")
    out = (C.c_int64 * (len(b) + 2))()
    n = fn(b, len(b), out, len(b) + 2)
    return [out[i] for i in range(n)]


def word_boundaries(text: str):
    return _bounds(lib().tbo_word_boundaries, text)


def sentence_boundaries(text: str):
    return _bounds(lib().tbo_sentence_boundaries, text)


def _pack(items):
    enc = [s.encode("utf-8") for s in items]
    off = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        off[1:] = np.cumsum([len(e) for e in enc])
    return b"".join(enc), off


def find_duplicates(items):
    blob, off = _pack(items)
    a, b = C.c_int64(), C.c_int64()
    lib().tbo_find_duplicates(blob, off.ctypes.data, len(items), C.byref(a), C.byref(b))
    return a.value, b.value


def find_top_duplicate(items):
    blob, off = _pack(items)
    return lib().tbo_find_top_duplicate(blob, off.ctypes.data, len(items))


def find_all_duplicate(words, n):
    blob, off = _pack(words)
    return lib().tbo_find_all_duplicate(blob, off.ctypes.data, len(words), n)


def get_n_grams(words, n):
    blob, off = _pack(words)
    cap = (len(blob) + len(words) + 2) * max(n, 1) + 16
    out = C.create_string_buffer(cap)
    cnt = lib().tbo_get_n_grams(blob, off.ctypes.data, len(words), n, out, cap)
    return out.value.decode("utf-8").split("\n") if cnt else []


def punctuation_contains(ch: str) -> bool:
    return bool(lib().tbo_punctuation_contains(ord(ch)))


def cinfo(cp: int) -> int:
    return lib().tbo_cinfo(cp)


def to_lowercase(text: str) -> str:
    b = text.encode("utf-8")
    out = C.create_string_buffer(3 * len(b) + 8)
    n = lib().tbo_to_lowercase(b, len(b), out, len(out))
    return out.raw[:n].decode("utf-8")


def decode_html_entities(text: str) -> str:
    """html_escape::decode_html_entities as restated in tb_oracle.cpp (parity unpinned)"""
    b = text.encode("utf-8")
    out = C.create_string_buffer(2 * len(b) + 8)
    n = lib().tbo_decode_html_entities(b, len(b), out, len(out))
    return out.raw[:n].decode("utf-8")


def remove_citations(text: str) -> str:
    b = text.encode("utf-8")
    out = C.create_string_buffer(len(b) + 8)
    n = lib().tbo_remove_citations(b, len(b), out, len(out))
    return out.raw[:n].decode("utf-8")


class OracleResult:
    def __init__(self, handle, steps_types, n):
        self._h = handle
        self.n = n
        L = lib()
        self.keep = np.ctypeslib.as_array(C.cast(L.tbo_result_keep(handle), C.POINTER(C.c_uint8)), (n,)).copy() if n else np.zeros(0, np.uint8)
        self.failed_step = np.ctypeslib.as_array(C.cast(L.tbo_result_failed_step(handle), C.POINTER(C.c_int8)), (n,)).copy() if n else np.zeros(0, np.int8)
        self.stats = []
        for k, t in enumerate(steps_types):
            dt = _abi.stats_dtype(t)
            if n:
                raw = C.string_at(L.tbo_result_stats(handle, k), n * dt.itemsize)
                self.stats.append(np.frombuffer(raw, dtype=dt).copy())
            else:
                self.stats.append(np.zeros(0, dt))

    def reason(self, i):
        return lib().tbo_result_reason(self._h, i).decode("utf-8")

    def metadata(self, i):
        md = lib().tbo_result_metadata(self._h, i).decode("utf-8")
        return dict(l.split("=", 1) for l in md.split("\n") if l)

    def content(self, i):
        ln = C.c_int64()
        p = lib().tbo_result_content(self._h, i, C.byref(ln))
        return C.string_at(p, ln.value).decode("utf-8")

    def content_packed(self):
        tot = lib().tbo_result_content_total(self._h)
        buf = np.zeros(max(tot, 1), np.uint8)
        off = np.zeros(self.n + 1, np.int64)
        lib().tbo_result_content_packed(self._h, buf.ctypes.data, off.ctypes.data)
        return buf[:tot], off

    def __del__(self):
        try:
            lib().tbo_result_free(self._h)
        except Exception:
            pass


class OraclePipeline:
    """steps: list of textblaster_b200.config param objects (C4QualityParams, ...)."""

    def __init__(self, steps):
        self._keep = []
        arr = (_abi.tb_step_cfg * max(len(steps), 1))()
        self.types = []
        for i, s in enumerate(steps):
            arr[i] = s.to_c(self._keep)
            self.types.append(arr[i].type)
        self._arr = arr
        self._h = lib().tbo_pipeline_create(C.addressof(arr), len(steps))
        if not self._h:
            raise RuntimeError("oracle: bad step config")

    def run_packed(self, utf8: np.ndarray, offsets: np.ndarray, nthreads: int = 1) -> OracleResult:
        utf8 = np.ascontiguousarray(utf8, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        n = len(offsets) - 1
        h = lib().tbo_run_batch(self._h, utf8.ctypes.data if utf8.size else None, offsets.ctypes.data, n, nthreads)
        return OracleResult(h, self.types, n)

    def run_texts(self, texts, nthreads: int = 1) -> OracleResult:
        blob, off = _pack(texts)
        return self.run_packed(np.frombuffer(blob, dtype=np.uint8), off, nthreads)

    def __del__(self):
        try:
            lib().tbo_pipeline_destroy(self._h)
        except Exception:
            pass
