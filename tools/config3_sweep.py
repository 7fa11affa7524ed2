"""BASELINE.json config 3: GopherRepetitionFilter alone on 20 000 long documents (10-200 KB, uniform in
the logarithm), one n-gram order at a time, n = 2..10, for the top-n-gram list (fraction 0.2) and the duplicate
n-gram list (fraction 0.15).  Reports documents/s per order and the share of documents whose tables do not fit shared
memory (tier X: more than 8192 words or line pieces, global work area).  Needs a GPU:
    python tools/config3_sweep.py > gpurun_out/r01_config3_sweep.json
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from textblaster_b200 import synth  # noqa: E402
from textblaster_b200.config import GopherRepetitionParams  # noqa: E402
from textblaster_b200.executor import GpuBatchPipeline  # noqa: E402

N_DOCS, LO, HI, SEED = 20000, 10_000, 200_000, synth.BASE_SEED + 3


def corpus():
    """Long documents made of consecutive synthetic C4-shaped ones: the offsets are a subset of the generator's."""
    rng = np.random.default_rng(SEED)
    targets = (LO * (HI / LO) ** rng.random(N_DOCS)).astype(np.int64)
    import torch
    holder = {}

    def pinned(nbytes):  # like bench.py: the text column in pinned host memory (pageable memory is copied at a fraction of the link rate)
        holder["t"] = torch.empty(max(nbytes, 1), dtype=torch.uint8, pin_memory=True)
        return holder["t"].numpy()

    utf8, off = synth.generate(int(targets.sum() / 3300) + 20000, seed=SEED, out=pinned)
    corpus.keep = holder  # the tensor owns the memory
    cut = np.searchsorted(off, np.minimum(np.cumsum(targets), off[-1]))
    cut = np.unique(np.concatenate([[0], cut]))
    return utf8, off[cut]


def main():
    utf8, off = corpus()
    n = len(off) - 1
    out = {"workload": "GopherRepetitionFilter n-gram sweep, %d documents of 10-200 KB (%.2f GB), end to end from pinned host memory" % (n, off[-1] / 1e9), "rows": []}
    for kind in ("top_n_grams", "dup_n_grams"):
        for order in range(2, 11):
            p = GpuBatchPipeline([GopherRepetitionParams(**{kind: [(order, 0.2 if kind == "top_n_grams" else 0.15)]})])
            p.run_packed(utf8, off)
            t0 = time.perf_counter()
            for _ in range(2):
                r = p.run_packed(utf8, off)
            dt = (time.perf_counter() - t0) / 2
            kt = {k: round(v[0] / 3, 3) for k, v in p.kernel_times().items()}
            st = r.stats[0]
            big = np.maximum(st["n_words"], st["n_lines"])
            out["rows"].append({"list": kind, "n": order, "docs_per_s": n / dt, "text_gbs": off[-1] / dt / 1e9, "kept": int(r.keep.sum()), "kernel_ms": kt,
                                "tier_share": {"S": float((big <= 512).mean()), "M": float(((big > 512) & (big <= 2048)).mean()),
                                               "L": float(((big > 2048) & (big <= 8192)).mean()), "X_global": float((big > 8192).mean())}})
            p.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
