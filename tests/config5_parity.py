"""BASELINE.json config 5 at full size: 200 000 multilingual documents (~2 KB each: Cyrillic, Greek,
Turkish, German, Danish, Hangul, Arabic, emoji / ZWJ / flags, mixed white space, combining marks, and the Han / Kana /
Thai subset that is reported separately) through the CUDA chain and through the oracle; every decision, reason bit
and integer counter has to be identical.  Test infrastructure (it runs the oracle); needs a GPU:
    python tests/config5_parity.py [n_docs] > gpurun_out/r01_config5_parity.json
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, HERE)
import multilingual  # noqa: E402
import oracle  # noqa: E402
from parity_cases import CHAINS, step_types  # noqa: E402
from parity_util import executed  # noqa: E402
from textblaster_b200 import _abi  # noqa: E402
from textblaster_b200.executor import GpuBatchPipeline  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    t0 = time.time()
    texts = multilingual.corpus(5, n)
    enc = [t.encode("utf-8") for t in texts]
    off = np.zeros(n + 1, np.int64)
    off[1:] = np.cumsum([len(e) for e in enc])
    utf8 = np.frombuffer(b"".join(enc), dtype=np.uint8)
    out = {"workload": "config 5: %d multilingual documents, %.1f MB" % (n, off[-1] / 1e6), "generator_s": round(time.time() - t0, 1), "chains": {}}
    for name in ("shipped_order", "baseline_order", "c4_then_open"):
        steps = CHAINS[name]
        p = GpuBatchPipeline(steps)
        t0 = time.time()
        g = p.run_packed(utf8, off)
        t_gpu = time.time() - t0
        t0 = time.time()
        o = oracle.OraclePipeline(steps).run_packed(utf8, off, nthreads=os.cpu_count() or 1)
        t_cpu = time.time() - t0
        bad = int((g.keep != o.keep).sum() + (g.failed_step != o.failed_step).sum())
        fields = 0
        for s_i, t in enumerate(step_types(steps)):
            ran = executed(o.failed_step, s_i)
            for f in g.stats[s_i].dtype.names:
                fields += 1
                bad += int((g.stats[s_i][f][ran] != o.stats[s_i][f][ran]).sum())
        ob, oo = o.content_packed()
        content_equal = bool(g.content is None or (np.array_equal(g.content_offsets, oo) and np.array_equal(g.content, ob)))
        flagged = int((g.stats[0]["flags"] & _abi.TB_DOCFLAG_COMPLEX_SCRIPT != 0).sum()) if "flags" in g.stats[0].dtype.names else None
        out["chains"][name] = {"documents": n, "kept": int(g.keep.sum()), "mismatching_values": bad, "counter_fields_compared": fields,
                               "content_identical": content_equal, "complex_script_documents": flagged,
                               "gpu_s": round(t_gpu, 3), "oracle_s": round(t_cpu, 2), "oracle_threads": os.cpu_count()}
        p.close()
    print(json.dumps(out))
    sys.exit(0 if all(c["mismatching_values"] == 0 and c["content_identical"] for c in out["chains"].values()) else 1)


if __name__ == "__main__":
    main()
