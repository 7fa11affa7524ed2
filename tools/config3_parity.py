"""BASELINE.json config 3 at full size, checked: the 10-200 KB corpus of tools/config3_sweep.py through
GopherRepetitionFilter with all nine n-gram orders (2..10) in both lists at once, on the GPU (C ABI) and through
the oracle on all host threads; every statistic and decision compared, tolerance 0.  Needs a GPU:
    python tools/config3_parity.py [--docs N] > gpurun_out/r02_config3_parity.json
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import config3_sweep  # noqa: E402
import oracle  # noqa: E402  (test infrastructure: the checker)
from parity_util import executed  # noqa: E402
from textblaster_b200.config import GopherRepetitionParams  # noqa: E402
from textblaster_b200.executor import GpuBatchPipeline  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--docs", type=int, default=0, help="check only the first N documents (0: all)")
    args = ap.parse_args()
    utf8, off = config3_sweep.corpus()
    if args.docs:
        off = off[: args.docs + 1]
    n = len(off) - 1
    orders = list(range(2, 11))
    params = GopherRepetitionParams(dup_line_frac=0.3, dup_para_frac=0.3, dup_line_char_frac=0.2, dup_para_char_frac=0.2,
                                    top_n_grams=[(k, 0.2) for k in orders], dup_n_grams=[(k, 0.15) for k in orders])
    t0 = time.perf_counter()
    res = GpuBatchPipeline([params]).run_packed(utf8, off)
    t_gpu = time.perf_counter() - t0
    threads = os.cpu_count() or 1
    t0 = time.perf_counter()
    ores = oracle.OraclePipeline([params]).run_packed(utf8, off.astype(np.int64), nthreads=threads)
    t_cpu = time.perf_counter() - t0
    mism = int(((res.keep != ores.keep) | (res.failed_step != ores.failed_step)).sum())
    values = 2 * n
    fields = {}
    ran = executed(ores.failed_step, 0)
    for name in res.stats[0].dtype.names:
        ne = res.stats[0][name][ran] != ores.stats[0][name][ran]
        if ne.ndim > 1:
            ne = ne.any(axis=1)
        fields[name] = int(ne.sum())
        mism += fields[name]
        values += int(ran.sum())
    big = np.maximum(res.stats[0]["n_words"], res.stats[0]["n_lines"])
    print(json.dumps({"workload": "GopherRepetitionFilter, top_n_grams and dup_n_grams with n = 2..10 each, %d documents of 10-200 KB (%.2f GB)" % (n, (off[-1] - off[0]) / 1e9),
                      "docs": n, "values_compared": values, "mismatches": mism, "mismatches_by_field": fields, "kept": int(res.keep.sum()),
                      "tier_X_share": float((big > 8192).mean()), "gpu_s_first_call": round(t_gpu, 2),
                      "oracle_s": round(t_cpu, 2), "oracle_threads": threads}))
    sys.exit(1 if mism else 0)


if __name__ == "__main__":
    main()
