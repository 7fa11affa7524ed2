"""tb_gopher_repetition_noalloc: the repetition filter for callers that must not allocate.

Some embedders cannot let this library call cudaMalloc on their thread -- their device memory is
parcelled out once by a pool somebody else owns, or they sit on a path where the CUDA allocator
(which may synchronise the whole device) is off limits.  For them the library carries a fixed 24 MiB
work array inside its own image; this entry point runs the GopherRepetitionFilter kernels out of that
array on caller-provided device buffers.  The price is a hard batch limit (2048 documents, 512 KiB).

The test drives it through the C ABI twice in a row (the second batch is different and smaller, so
anything stale in the shared array would show), compares statistics and keep flags with the regular
allocating entry point and with the CPU oracle, and checks that device memory in use does not move
across the call.  The very first call (the one that does the set-up) is made on a non-blocking stream:
its kernels must not start before the tables have landed.  Further tests fill the array to its limit
(2048 documents, 512 KiB) and hand over offsets that break the contract.
This file is named so that pytest collects it last.
"""
import ctypes as C
import os

import numpy as np
import pytest

import oracle
from parity_cases import GR_SHIPPED
from textblaster_b200 import _abi, _native, synth

pytestmark = pytest.mark.gpu


def _entry():
    lib = _native.lib()
    fn = lib.tb_gopher_repetition_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
    return fn


def _run_noalloc(fn, utf8, off, params_c, bufs, stream=None, nb=None):
    import torch

    d_text, d_off, d_stats, d_keep = bufs
    n = len(off) - 1
    nb = int(off[-1]) if nb is None else nb
    d_text[:nb].copy_(torch.from_numpy(utf8[:nb].copy()))
    d_off[:n + 1].copy_(torch.from_numpy(off.copy()))
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    err = C.create_string_buffer(512)
    rc = fn(d_text.data_ptr(), d_off.data_ptr(), n, nb, C.addressof(params_c), d_stats.data_ptr(), d_keep.data_ptr(),
            C.c_void_p(stream.cuda_stream) if stream is not None else None, err, len(err))
    assert rc == _abi.TB_OK, err.value
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    dt = _abi.stats_dtype(_abi.TB_STEP_GOPHER_REPETITION)
    stats = np.frombuffer(d_stats[: n * dt.itemsize].cpu().numpy().tobytes(), dtype=dt)
    keep = d_keep[:n].cpu().numpy()
    return stats, keep, free_before, free_after


def test_noalloc_matches_the_allocating_entry_point():
    import torch

    from textblaster_b200.executor import GpuBatchPipeline

    fn = _entry()
    step = GR_SHIPPED.to_c([])
    params_c = step.u.gr
    dt = _abi.stats_dtype(_abi.TB_STEP_GOPHER_REPETITION)
    # caller-owned device buffers, allocated up front (the entry point itself must not allocate)
    bufs = (torch.empty(512 * 1024 + 64, dtype=torch.uint8, device="cuda"), torch.empty(2049, dtype=torch.int64, device="cuda"),
            torch.empty(2048 * dt.itemsize, dtype=torch.uint8, device="cuda"), torch.empty(2048, dtype=torch.uint8, device="cuda"))
    ref = GpuBatchPipeline([GR_SHIPPED])
    batches = [synth.generate(220, seed=41, median_bytes=1500), synth.generate(90, seed=42, median_bytes=900)]
    side = torch.cuda.Stream()  # torch creates its streams with cudaStreamNonBlocking
    orc = oracle.OraclePipeline([GR_SHIPPED])
    for i, (utf8, off) in enumerate(batches):
        assert int(off[-1]) <= 512 * 1024
        stats, keep, free_before, free_after = _run_noalloc(fn, utf8, off, params_c, bufs, stream=side if i == 0 else None)
        want = ref.run_packed(utf8, off)
        assert np.array_equal(keep, want.keep), f"call {i}: keep flags differ"
        for name in dt.names:
            assert np.array_equal(stats[name], want.stats[0][name]), f"call {i}: {name} differs"
        # and directly against the CPU oracle (tolerance 0: integer counters)
        ores = orc.run_packed(utf8, off.astype(np.int64), nthreads=os.cpu_count() or 1)
        assert np.array_equal(keep, ores.keep), f"call {i}: keep flags differ from the oracle"
        for name in dt.names:
            assert np.array_equal(stats[name], ores.stats[0][name]), f"call {i}: {name} differs from the oracle"
        if i > 0:  # the first call may do its one-time set-up; afterwards nothing may be allocated
            assert free_after == free_before, "device memory in use changed across a no-alloc call"
    assert 0 < int(want.keep.sum()) <= len(want.keep)
    ref.close()


def test_noalloc_rejects_oversized_batches():
    fn = _entry()
    step = GR_SHIPPED.to_c([])
    err = C.create_string_buffer(512)
    one = C.c_int64(0)
    rc = fn(C.addressof(one), C.addressof(one), 2049, 10, C.addressof(step.u.gr), C.addressof(one), C.addressof(one), None, err, len(err))
    assert rc == _abi.TB_ERR_INVALID_ARGUMENT and b"2048 documents" in err.value
    rc = fn(C.addressof(one), C.addressof(one), 10, 512 * 1024 + 1, C.addressof(step.u.gr), C.addressof(one), C.addressof(one), None, err, len(err))
    assert rc == _abi.TB_ERR_INVALID_ARGUMENT and b"524288 bytes" in err.value


def _buffers():
    import torch

    dt = _abi.stats_dtype(_abi.TB_STEP_GOPHER_REPETITION)
    return (torch.empty(512 * 1024 + 64, dtype=torch.uint8, device="cuda"), torch.empty(2049, dtype=torch.int64, device="cuda"),
            torch.empty(2048 * dt.itemsize, dtype=torch.uint8, device="cuda"), torch.empty(2048, dtype=torch.uint8, device="cuda"))


def test_noalloc_at_the_limit():
    """exactly TB_NOALLOC_MAX_DOCS documents and TB_NOALLOC_MAX_BYTES bytes, dense in words: the carve at capacity"""
    fn = _entry()
    step = GR_SHIPPED.to_c([])
    dt = _abi.stats_dtype(_abi.TB_STEP_GOPHER_REPETITION)
    raw, roff = synth.generate(2047, seed=43, median_bytes=120)
    parts = []
    for d in range(2047):  # short documents: each cut at a space (never inside a character) to at most 150 bytes
        doc = raw[int(roff[d]):int(roff[d + 1])].tobytes()
        parts.append(doc if len(doc) <= 150 else doc[: doc.rfind(b" ", 0, 150)])
    utf8 = np.frombuffer(b"".join(parts), np.uint8)
    off = np.zeros(2048, np.int64)
    off[1:] = np.cumsum([len(p) for p in parts])
    assert int(off[-1]) < 400 * 1024
    # the last document fills the text up to the limit with the densest word pattern there is (two words per
    # three bytes), which also makes it a tier-X document
    fill = 512 * 1024 - int(off[-1])
    last = (("a\u00a9" * (fill // 3 + 1)).encode("utf-8"))[: fill - fill % 3] + b"x" * (fill % 3)
    assert len(last) == fill
    utf8 = np.concatenate([utf8[: int(off[-1])], np.frombuffer(last, np.uint8)])
    off = np.concatenate([off, [512 * 1024]]).astype(np.int64)
    assert len(off) - 1 == 2048 and int(off[-1]) == 512 * 1024
    stats, keep, _, _ = _run_noalloc(fn, utf8, off, step.u.gr, _buffers())
    ores = oracle.OraclePipeline([GR_SHIPPED]).run_packed(utf8, off, nthreads=os.cpu_count() or 1)
    assert np.array_equal(keep, ores.keep)
    for name in dt.names:
        assert np.array_equal(stats[name], ores.stats[0][name]), f"{name} differs from the oracle"
    assert int(stats["n_words"][-1]) > 100000


def test_noalloc_skips_documents_with_bad_offsets():
    """offsets live in device memory and cannot be checked on the host without a synchronisation: a document whose
    offsets decrease or leave [0, n_bytes] is skipped on the device and flagged, its neighbours are unaffected"""
    fn = _entry()
    step = GR_SHIPPED.to_c([])
    dt = _abi.stats_dtype(_abi.TB_STEP_GOPHER_REPETITION)
    utf8, off = synth.generate(40, seed=44, median_bytes=600)
    off = off.astype(np.int64)
    good = oracle.OraclePipeline([GR_SHIPPED]).run_packed(utf8, off, nthreads=1)
    bad = off.copy()
    bad[11] = bad[10] - 5          # document 10 ends before it starts; document 11 starts inside document 9
    bad[21] = int(off[-1]) + 4096  # document 20 ends beyond n_bytes; 21 starts there
    bad[31] = -7                   # 30 decreasing, 31 negative start
    stats, keep, _, _ = _run_noalloc(fn, utf8, bad, step.u.gr, _buffers(), nb=int(off[-1]))
    broken = {10, 11, 20, 21, 30, 31}
    for d in range(40):
        if d in broken:
            assert keep[d] == 0 and stats["flags"][d] & _abi.TB_DOCFLAG_BAD_OFFSETS and stats["n_words"][d] == 0
        else:
            assert keep[d] == good.keep[d]
            for name in dt.names:
                assert np.array_equal(stats[name][d], good.stats[0][name][d]), f"doc {d}: {name}"
