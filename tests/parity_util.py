"""Shared helpers for parity tests: compare a result (host emulation or the CUDA library through the
C ABI) with the oracle, field by field, bit-exact."""
import numpy as np

from textblaster_b200 import _abi


def executed(failed_step, s):
    """documents for which step s actually ran (executor short-circuit, src/executor.rs:32-46)"""
    fs = np.asarray(failed_step).astype(np.int64)
    return (fs < 0) | (fs >= s)


def compare(oracle_res, keep, failed_step, stats, contents, types, texts=None, label=""):
    """Raises AssertionError with the first mismatching document."""
    n = len(keep)
    okeep, ofail = oracle_res.keep, oracle_res.failed_step

    def show(i):
        t = texts[i] if texts is not None else None
        return f"{label} doc {i}: {t!r:.300}"

    for i in range(n):
        assert int(keep[i]) == int(okeep[i]) and int(failed_step[i]) == int(ofail[i]), (
            f"{show(i)} keep/failed_step: got ({keep[i]}, {failed_step[i]}) oracle ({okeep[i]}, {ofail[i]})"
            f" oracle reason {oracle_res.reason(i)!r}")
    for s, t in enumerate(types):
        ran = executed(ofail, s)
        a, b = stats[s], oracle_res.stats[s]
        for name in a.dtype.names:
            x, y = a[name][ran], b[name][ran]
            if not np.array_equal(x, y):
                idx = np.nonzero(ran)[0]
                bad = [j for j in range(len(idx)) if not np.array_equal(x[j], y[j])][0]
                raise AssertionError(f"{show(int(idx[bad]))} step {s} ({_abi.STEP_NAMES[t]}) field {name}: got {x[bad]} oracle {y[bad]}")
    if contents is not None:
        for i in range(n):
            oc = oracle_res.content(i)
            assert contents[i] == oc, f"{show(i)} content: got {contents[i]!r:.300} oracle {oc!r:.300}"
