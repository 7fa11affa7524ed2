"""Cross-check helper: ICU4C 71 break iterators (bundled with Nsight Compute) via ctypes.

NOT the reference (ICU4X 1.5) -- a proxy used only to fuzz the oracle's UAX #29
restatement on rule-based scripts. Absent library => callers skip.
"""
import ctypes, glob, os

_UBRK_WORD, _UBRK_SENTENCE = 1, 3
_lib = None


def _load():
    global _lib
    if _lib is not None:
        return _lib
    pats = glob.glob("/opt/nvidia/nsight-compute/*/host/linux-desktop-glibc_2_11_3-x64/")
    for d in pats:
        try:
            ctypes.CDLL(os.path.join(d, "libicudata.so.71"), mode=ctypes.RTLD_GLOBAL)
            uc = ctypes.CDLL(os.path.join(d, "libicuuc.so.71"), mode=ctypes.RTLD_GLOBAL)
            _lib = uc
            return uc
        except OSError:
            continue
    _lib = False
    return False


def available():
    return bool(_load())


def breaks(text: str, kind: str = "word", locale: bytes = b"root"):
    """Return break offsets in UTF-16 code units converted to code point indices."""
    lib = _load()
    if not lib:
        raise RuntimeError("ICU4C not found")
    u16 = text.encode("utf-16-le")
    n16 = len(u16) // 2
    buf = (ctypes.c_uint16 * (n16 + 1)).from_buffer_copy(u16 + b"\0\0")
    err = ctypes.c_int(0)
    lib.ubrk_open_71.restype = ctypes.c_void_p
    lib.ubrk_open_71.argtypes = [ctypes.c_int, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_int)]
    bi = lib.ubrk_open_71(_UBRK_WORD if kind == "word" else _UBRK_SENTENCE, locale, buf, n16, ctypes.byref(err))
    if err.value > 0:
        raise RuntimeError(f"ubrk_open failed {err.value}")
    lib.ubrk_next_71.argtypes = [ctypes.c_void_p]
    lib.ubrk_next_71.restype = ctypes.c_int32
    lib.ubrk_close_71.argtypes = [ctypes.c_void_p]
    out16 = [0]
    while True:
        p = lib.ubrk_next_71(bi)
        if p < 0:
            break
        out16.append(p)
    lib.ubrk_close_71(bi)
    # map utf16 offsets -> code point offsets
    cp_of_u16 = {}
    k = 0
    for i, ch in enumerate(text):
        cp_of_u16[k] = i
        k += 2 if ord(ch) > 0xFFFF else 1
    cp_of_u16[k] = len(text)
    return [cp_of_u16[o] for o in out16]


def segments(text, kind="word"):
    b = breaks(text, kind)
    return [text[b[i]:b[i + 1]] for i in range(len(b) - 1)]


if __name__ == "__main__":
    import sys
    for t in sys.argv[1:] or ["a:b", "mid...dle", "can't 3.14 foo_bar a.b 1,000", "Hello. How are you? Fine! ok"]:
        print(repr(t), segments(t, "word"), segments(t, "sentence"))
