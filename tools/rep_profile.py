"""Where the repetition pass spends a document's time, by tier and phase: needs a measurement build of the library
(tools/build_variant.sh prof -DTB_REP_PROFILE) loaded through TB_LIB_PATH.  Clocks are thread 0's, per document."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np  # noqa: E402

import bench  # noqa: E402
from textblaster_b200 import _abi, _native, synth  # noqa: E402
from textblaster_b200.executor import GpuBatchPipeline  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
utf8, off = synth.generate(n, seed=synth.BASE_SEED + 1)
pipe = GpuBatchPipeline(bench.shipped_chain())
pipe.set_option(_abi.TB_OPT_RETURN_CONTENT, 2)
lib = _native.lib()
buf = (C.c_ulonglong * 256)()
pipe.run_packed(utf8, off)
assert lib.tb_debug_rep_profile(buf) == 0  # (clears)
pipe.run_packed(utf8, off)
assert lib.tb_debug_rep_profile(buf) == 0
p = np.array(list(buf), np.float64).reshape(8, 32)
names = {1: "lines/paras", 2: "word offsets", 12: "top1: keys", 13: "top1: insert", 14: "top1: best", 7: "top1: compact", 3: "top: rest", 15: "sig: shortest", 16: "sig: clear", 17: "sig: insert", 4: "sig: compact", 5: "dup: levels", 6: "result"}
ORDER = (1, 2, 12, 13, 14, 7, 3, 15, 16, 17, 4, 5, 6)
tiers = ["w256", "w512", "c1k", "c2k", "c4k", "c8k", "X"]
print("%-6s %8s %9s | " % ("tier", "docs", "us/doc") + " ".join("%13s" % names[k] for k in ORDER) + " | items<=32 <=256 more")
for t in range(7):
    docs = p[t][8]
    if not docs:
        continue
    tot = sum(p[t][k] for k in names)
    print("%-6s %8d %9.1f | " % (tiers[t], docs, tot / docs / 1.9e3) + " ".join("%6.1f (%3.0f%%)" % (p[t][k] / docs / 1.9e3, 100 * p[t][k] / tot) for k in ORDER)
          + " | %d %d %d" % (p[t][9], p[t][10], p[t][11]))
