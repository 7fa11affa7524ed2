"""A compact run of the table / register paths of the repetition pass and of the C4 bitmap passes, meant to be run under
compute-sanitizer (memcheck, racecheck) on a GPU box:
    compute-sanitizer --tool racecheck python tests/sanitizer_target.py
It checks its results against the oracle as well (the sanitizer slows the kernels down 10-100x: the corpus is small)."""
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

import fuzzgen  # noqa: E402
import oracle  # noqa: E402
from parity_cases import CHAINS, step_types  # noqa: E402
from parity_util import compare  # noqa: E402
from test_emul_parity import signature_pruning_docs, split_ambiguity_docs  # noqa: E402
from textblaster_b200 import _abi, synth  # noqa: E402
from textblaster_b200.executor import GpuBatchPipeline  # noqa: E402


def main():
    rng = random.Random(3)
    base = fuzzgen.english_doc(rng, 1500)
    texts = ([fuzzgen.english_doc(rng, sz) for sz in (300, 2500, 12000, 40000)] + [(base + "\n") * 4, "a b c d e f g " * 400, "x y " * 3000]
             + signature_pruning_docs()[:24] + split_ambiguity_docs()[:12] + fuzzgen.dense_word_docs()[:6] + fuzzgen.FuzzGen(5).docs(60))
    utf8, off = synth.generate(400, seed=9)
    texts += [bytes(utf8[off[i]:off[i + 1]]).decode("utf-8") for i in range(400)]
    for name, mode in (("gr_open", 1), ("shipped_order", 2), ("baseline_order", 1)):
        steps = CHAINS[name]
        p = GpuBatchPipeline(steps)
        p.set_option(_abi.TB_OPT_RETURN_CONTENT, mode)
        res = p.run_texts(texts)
        ores = oracle.OraclePipeline(steps).run_texts(texts, nthreads=os.cpu_count() or 1)
        contents = [res.content_of(i) for i in range(res.n_docs)] if res.content is not None else None
        compare(ores, res.keep, res.failed_step, res.stats, contents, step_types(steps), texts, label=name)
        p.close()
        print(name, "ok:", res.n_docs, "documents,", int(res.keep.sum()), "kept")


if __name__ == "__main__":
    main()
