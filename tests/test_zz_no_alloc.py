"""tb_gopher_repetition_noalloc: the repetition filter for callers that must not allocate.

Some embedders cannot let this library call cudaMalloc on their thread -- their device memory is
parcelled out once by a pool somebody else owns, or they sit on a path where the CUDA allocator
(which may synchronise the whole device) is off limits.  For them the library carries a fixed 16 MiB
work array inside its own image; this entry point runs the GopherRepetitionFilter kernels out of that
array on caller-provided device buffers.  The price is a hard batch limit (2048 documents, 512 KiB).

The test drives it through the C ABI twice in a row (the second batch is different and smaller, so
anything stale in the shared array would show), compares statistics and keep flags with the regular
allocating entry point, and checks that device memory in use does not move across the call.
This file is named so that pytest collects it last.
"""
import ctypes as C

import numpy as np
import pytest

from parity_cases import GR_SHIPPED
from textblaster_b200 import _abi, _native, synth

pytestmark = pytest.mark.gpu


def _entry():
    lib = _native.lib()
    fn = lib.tb_gopher_repetition_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
    return fn


def _run_noalloc(fn, utf8, off, params_c, bufs):
    import torch

    d_text, d_off, d_stats, d_keep = bufs
    n, nb = len(off) - 1, int(off[-1])
    d_text[:nb].copy_(torch.from_numpy(utf8[:nb].copy()))
    d_off[:n + 1].copy_(torch.from_numpy(off.copy()))
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    err = C.create_string_buffer(512)
    rc = fn(d_text.data_ptr(), d_off.data_ptr(), n, nb, C.addressof(params_c), d_stats.data_ptr(), d_keep.data_ptr(), None, err, len(err))
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
    for i, (utf8, off) in enumerate(batches):
        assert int(off[-1]) <= 512 * 1024
        stats, keep, free_before, free_after = _run_noalloc(fn, utf8, off, params_c, bufs)
        want = ref.run_packed(utf8, off)
        assert np.array_equal(keep, want.keep), f"call {i}: keep flags differ"
        for name in dt.names:
            assert np.array_equal(stats[name], want.stats[0][name]), f"call {i}: {name} differs"
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
