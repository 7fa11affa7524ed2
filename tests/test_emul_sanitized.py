"""The kernel bodies under AddressSanitizer and UndefinedBehaviorSanitizer.

compute-sanitizer is not available on the GPU pool this was developed on, so the memory-safety evidence for the kernel
bodies comes from the host emulation: the same source (csrc/*.cuh), compiled with -fsanitize=address,undefined, run on
documents that reach every tier and both word paths, with every scratch array allocated at exactly the size the product
gives it (an index past an array's end is a heap-buffer-overflow here).  The child process also compares the results
with the oracle.  What this cannot see: races between lanes (the emulation runs the lanes of a warp as fibers of one
thread) -- the GPU tests cover those by repetition (test_gpu_parity.py::test_repeated_runs_are_identical)."""
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))

CHILD = r"""
import random, sys
import _emul, fuzzgen, oracle
from parity_cases import CHAINS, step_types
from parity_util import compare
from test_emul_parity import signature_pruning_docs, split_ambiguity_docs
rng = random.Random(5)
texts = [fuzzgen.english_doc(rng, sz) for sz in (150, 1200, 2600, 5200, 11000, 22000, 45000, 70000)]   # one document per tier of the repetition pass
texts += ["a b c d e f g " * 300, "x y " * 1500, ("the same line again\n" * 40), "p one.\n\np two.\n\n" * 30, "", " ", "\n\n"]
texts += signature_pruning_docs()[:10] + split_ambiguity_docs()[:8] + fuzzgen.dense_word_docs()[:4] + fuzzgen.FuzzGen(9).docs(40)
for name in ("baseline_order", "shipped_order", "gr_open", "gq_open", "c4_lax", "fw_open"):
    steps = CHAINS[name]
    o = oracle.OraclePipeline(steps).run_texts(texts)
    e = _emul.run_chain(steps, texts)
    compare(o, e.keep, e.failed_step, e.stats, e.contents, step_types(steps), texts, label=name)
    print(name, "ok", len(texts), flush=True)
"""


def _runtime(name):
    p = subprocess.run(["gcc", "-print-file-name=" + name], capture_output=True, text=True).stdout.strip()
    return p if os.path.isabs(p) and os.path.exists(p) else None


@pytest.mark.timeout(900)
def test_kernel_bodies_under_asan_ubsan():
    asan, ubsan = _runtime("libasan.so"), _runtime("libubsan.so")
    if not asan or not ubsan:
        pytest.skip("sanitizer runtimes not installed")
    env = dict(os.environ, TB_EMUL_SANITIZE="1", LD_PRELOAD=asan + ":" + ubsan, PYTHONPATH=HERE + os.pathsep + os.path.dirname(HERE),
               # leaks: CPython's own; fibers switch stacks under ASan's feet, so no stack-use-after-return mode
               ASAN_OPTIONS="detect_leaks=0:detect_stack_use_after_return=0:abort_on_error=0:halt_on_error=1",
               UBSAN_OPTIONS="print_stacktrace=1:halt_on_error=1")
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, cwd=HERE)
    log = r.stdout + r.stderr
    assert r.returncode == 0, log[-4000:]
    assert "AddressSanitizer" not in log and "runtime error" not in log, log[-4000:]
    assert log.count(" ok ") == 6
