"""tb_gather_content is pure host code (no CUDA call): the content of selected documents rebuilt from runs of the input
column.  Checked here without a GPU against a direct Python reconstruction -- random runs, a selection mask, several
thread counts, i32 and i64 offsets, the sizes-only call and the capacity check."""
import ctypes as C

import numpy as np
import pytest

from textblaster_b200 import _abi, _native
from textblaster_b200.executor import gather_content


def _case(seed, n, i32=False):
    rng = np.random.default_rng(seed)
    lens = rng.integers(0, 3000, n)
    off = np.zeros(n + 1, np.int64)
    off[1:] = np.cumsum(lens)
    utf8 = rng.integers(0, 256, max(int(off[-1]), 1), dtype=np.uint8)
    runs, ro = [], [0]
    for d in range(n):
        L, p = int(lens[d]), 0
        while p < L:
            b = p + int(rng.integers(0, 40))
            if b >= L:
                break
            e = min(L, b + int(rng.integers(1, 900)))
            runs.append((b, e))
            p = e + 1
        ro.append(len(runs))
    runs = np.array(runs, np.uint32).reshape(-1, 2)
    ro = np.array(ro, np.int64)
    out = _abi.tb_batch_out()
    out.n_docs, out.content_kind, out.n_runs = n, 2, len(runs)
    out.runs = runs.ctypes.data_as(C.POINTER(_abi.tb_run))
    out.run_offsets = ro.ctypes.data_as(C.POINTER(C.c_int64))
    return out, utf8, (off.astype(np.int32) if i32 else off), runs, ro, rng


def _expected(utf8, off, runs, ro, sel):
    parts, wo = [], [0]
    for d in range(len(off) - 1):
        if sel is not None and not sel[d]:
            continue
        r = runs[ro[d]:ro[d + 1]].astype(np.int64)
        parts += [utf8[int(off[d]) + b:int(off[d]) + e] for b, e in r]
        wo.append(wo[-1] + int((r[:, 1] - r[:, 0]).sum()))
    return (np.concatenate(parts) if parts else np.zeros(0, np.uint8)), np.array(wo, np.int64)


@pytest.mark.parametrize("threads", [1, 3, 8])
@pytest.mark.parametrize("i32", [False, True])
def test_gather_matches_python(threads, i32):
    out, utf8, off, runs, ro, rng = _case(7 + threads, 6000, i32)
    for sel in (None, (rng.random(6000) < 0.8).astype(np.uint8), np.zeros(6000, np.uint8)):
        got, goff = gather_content(out, utf8, off, select=sel, nthreads=threads)
        want, woff = _expected(utf8, off, runs, ro, sel)
        assert np.array_equal(goff, woff) and np.array_equal(got, want)


def test_gather_sizes_only_and_capacity():
    out, utf8, off, runs, ro, _ = _case(3, 500)
    lib = _native.lib()
    doff = np.zeros(501, np.int64)
    total = C.c_int64()
    assert lib.tb_gather_content(C.byref(out), utf8.ctypes.data, off.ctypes.data, 1, None, None, 0, doff.ctypes.data, C.byref(total), 2) == 0
    want, woff = _expected(utf8, off, runs, ro, None)
    assert total.value == len(want) and np.array_equal(doff, woff)
    small = np.zeros(max(len(want) - 1, 1), np.uint8)
    rc = lib.tb_gather_content(C.byref(out), utf8.ctypes.data, off.ctypes.data, 1, None, small.ctypes.data, small.size, doff.ctypes.data,
                               C.byref(total), 2)
    assert rc == _abi.TB_ERR_INVALID_ARGUMENT and b"needed" in lib.tb_last_error(None)
    out.content_kind = 1
    assert lib.tb_gather_content(C.byref(out), utf8.ctypes.data, off.ctypes.data, 1, None, None, 0, doff.ctypes.data, C.byref(total), 1) == _abi.TB_ERR_INVALID_ARGUMENT
