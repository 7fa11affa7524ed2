#!/usr/bin/env python3
"""bench.py -- docs/s and text GB/s through the C4 + Gopher filter chain (BASELINE.json metric).

Workload (BASELINE.json configs[1]): C4QualityFilter + GopherQualityFilter + GopherRepetitionFilter
over 1,000,000 synthetic C4-shaped documents (~2 KB median, log-normal) per GPU.  One "step" is one
pass of the whole chain over that batch.  Two measurements per run:

  value : inputs already resident in HBM (tb_pipeline_run_batch_device), results copied back;
  e2e   : the call a user makes (tb_pipeline_run_batch) on pinned HOST buffers -- H2D of the text
          and offsets and D2H of keep mask / statistics / rewritten content inside the timed region.

N > 1 (torchrun): documents are sharded across ranks (weak scaling: 1M docs per GPU), no data-path
collective -- the path is embarrassingly parallel; NCCL is used only for the timing barrier / max.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--docs D] [--impl reference]
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

# stdout carries exactly one JSON line: anything libraries print there (NCCL's version banner, for one) is
# sent to stderr by pointing fd 1 at fd 2 for the duration of the run; emit() writes to the real stdout
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
_REAL_STDOUT = None


def guard_stdout():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(line) + "\n").encode())

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

METRIC = "docs/s through C4+Gopher filter chain"
UNIT = "docs/s"
WORKLOAD = ("C4QualityFilter+GopherQualityFilter+GopherRepetitionFilter chain, 1M synthetic C4-shaped docs (~2 KB median) "
            "per GPU (BASELINE.json configs[1])")
GROUP_DOCS = 65536  # row-group size of the synthetic corpus


def shipped_chain(with_fineweb=False, order="baseline"):
    """BASELINE.json configs[1] order (C4 -> GopherQuality -> GopherRepetition), parameters of the
    reference's config/pipeline_config.yaml:6-32,56-66 with the default English stop words."""
    from textblaster_b200.config import C4QualityParams, FineWebQualityParams, GopherQualityParams, GopherRepetitionParams

    c4 = C4QualityParams(True, True, True, 5, 3, 1000, True, True, True, True)
    gq = GopherQualityParams(min_doc_words=50, max_doc_words=100000, min_avg_word_length=3.0, max_avg_word_length=10.0,
                             max_symbol_word_ratio=0.1, max_bullet_lines_ratio=0.9, max_ellipsis_lines_ratio=0.3,
                             max_non_alpha_words_ratio=0.8, min_stop_words=2)
    gr = GopherRepetitionParams(dup_line_frac=0.3, dup_para_frac=0.3, dup_line_char_frac=0.2, dup_para_char_frac=0.2,
                                top_n_grams=[(2, 0.2), (3, 0.18), (4, 0.16)],
                                dup_n_grams=[(5, 0.15), (6, 0.14), (7, 0.13), (8, 0.12), (9, 0.11), (10, 0.10)])
    # order="shipped": the order of config/pipeline_config.yaml (GopherRepetition -> GopherQuality -> C4), SURVEY 8d config 4
    chain = [gr, gq, c4] if order == "shipped" else [c4, gq, gr]
    if with_fineweb:  # SURVEY 8(f) row 3, pipeline_config.yaml:74-80; not part of the headline workload
        chain.append(FineWebQualityParams(0.12, False, 0.67, 30, 0.01, 0.3))
    return chain


def pin_to_gpu_socket(torch, index):
    """Run this process (and so place its pinned host buffers, by first touch) on the CPU socket the GPU hangs off
    (one reader per GPU, NUMA-pinned).  Returns a short description for the config line."""
    try:
        pr = torch.cuda.get_device_properties(index)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        base = "/sys/bus/pci/devices/" + bdf
        node = int(open(base + "/numa_node").read().strip())
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            if part:
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if node < 0 or not cpus:
            return "not pinned (no NUMA information for %s)" % bdf
        os.sched_setaffinity(0, cpus)
        return "pinned to NUMA node %d (%d cpus)" % (node, len(cpus))
    except Exception as e:  # no sysfs, odd device properties: run unpinned
        return "not pinned (%s)" % type(e).__name__


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel, chunk_bytes_per_launch):
    """dram bytes per launch of the dominant kernel from the committed ncu capture, if any.  The capture is of ONE
    launch (the algorithmic bytes of the chunk it worked on are in the file); launches differ in size, so the figure is
    scaled by chunk size to the average launch that `achieved` is quoted on (a tier kernel of the repetition pass works
    on its share of every chunk: the same ratio applies)."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            t, cap = d.get(kernel), d.get("_captured_launch_algorithmic_bytes")
            if t is None:
                return None
            return t * chunk_bytes_per_launch / cap if cap else t
        except Exception:
            return None
    return None


def cpu_baseline(utf8, off, steps, budget_s=12.0, threads=None):
    """The oracle (C++ restatement of the reference CPU chain, kind "port") timed on this box's host
    cores on a bounded sample of the same workload."""
    import oracle

    threads = threads or (os.cpu_count() or 1)
    pl = oracle.OraclePipeline(steps)
    n = len(off) - 1
    probe = min(n, 2000)
    t0 = time.perf_counter()
    pl.run_packed(utf8[off[0]:off[probe]], off[:probe + 1] - off[0], nthreads=threads)
    rate = probe / max(time.perf_counter() - t0, 1e-6)
    sample = int(min(n, max(probe, rate * budget_s)))
    t0 = time.perf_counter()
    r = pl.run_packed(utf8[off[0]:off[sample]], off[:sample + 1] - off[0], nthreads=threads)
    dt = time.perf_counter() - t0
    nbytes = int(off[sample] - off[0])
    return {"value": sample / dt, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"first {sample} docs ({nbytes / 1e6:.1f} MB) of the workload, {dt:.2f} s, oracle/tb_oracle.cpp with {threads} threads",
            "text_gbs": nbytes / dt / 1e9, "keep_rate": float(r.keep.mean()) if sample else None}, r, sample


def parity_check(pipe, steps, utf8, off, sample, ores):
    """The benchmarked run, checked: the GPU's results for the documents of the cpu_baseline sample against the
    oracle's for the same documents -- keep, failing step, every statistics field of every step that ran, and the
    rewritten content byte for byte.  Tolerance 0.  Returns {"docs", "values", "mismatches", ...}."""
    from parity_util import executed

    sub_off = (off[:sample + 1] - off[0]).astype(np.int64)
    res = pipe.run_packed(utf8[off[0]:off[sample]], sub_off)
    mism, values, first = 0, 0, None

    def note(what, idx):
        nonlocal first
        if first is None and len(idx):
            first = f"{what} at doc {int(idx[0])}"

    bad = np.flatnonzero((res.keep != ores.keep) | (res.failed_step != ores.failed_step))
    mism += len(bad); values += 2 * sample
    note("keep/failed_step", bad)
    for s in range(len(steps)):
        ran = executed(ores.failed_step, s)
        a, b = res.stats[s], ores.stats[s]
        for name in a.dtype.names:
            x, y = a[name][ran], b[name][ran]
            ne = x != y
            if ne.ndim > 1:
                ne = ne.any(axis=1)
            bad = np.flatnonzero(ran)[ne]
            mism += len(bad); values += int(ran.sum())
            note(f"step {s} field {name}", bad)
    content = None
    if res.content is not None:
        ob, oo = ores.content_packed()
        same = np.array_equal(res.content_offsets, oo) and np.array_equal(res.content, ob)
        content = bool(same)
        values += sample
        if not same:
            lens_a, lens_b = np.diff(res.content_offsets), np.diff(oo)
            bad = np.flatnonzero(lens_a != lens_b)
            mism += max(len(bad), 1)
            note("content", bad if len(bad) else np.array([0]))
    out = {"docs": int(sample), "values": int(values), "mismatches": int(mism), "content_identical": content,
           "checked": "GPU (tb_pipeline_run_batch) vs oracle on the cpu_baseline sample: keep, failed_step, all statistics of executed steps, rewritten content; tolerance 0"}
    if first:
        out["first_mismatch"] = first
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation is Rust + ICU4X and cannot be built in
    this image (no cargo/rustc), so this arm times the oracle port on all host threads."""
    if rank != 0:
        return
    from textblaster_b200 import synth

    steps = shipped_chain(args.with_fineweb, args.order)
    threads = os.cpu_count() or 1
    import oracle

    pl = oracle.OraclePipeline(steps)
    probe_u, probe_o = synth.generate(2000, seed=synth.BASE_SEED + 2)
    t0 = time.perf_counter()
    pl.run_packed(probe_u, probe_o, nthreads=threads)
    rate = 2000 / max(time.perf_counter() - t0, 1e-6)
    total_steps = args.steps + args.warmup
    sample = int(max(2000, min(args.docs, rate * min(20.0, 150.0 / max(total_steps, 1)))))
    utf8, off = synth.generate(sample, seed=synth.BASE_SEED + 2)
    for _ in range(args.warmup):
        pl.run_packed(utf8, off, nthreads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = pl.run_packed(utf8, off, nthreads=threads)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    v = sample / dt
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD + (" + FineWebQualityFilter" if args.with_fineweb else "") + (" [shipped step order]" if args.order == "shipped" else ""), "docs_per_step": sample, "bytes_per_step": int(off[-1]),
                       "note": "CPU arm: each step is a bounded sample of the workload"},
            "text_gbs": float(off[-1]) / dt / 1e9,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{sample} docs ({off[-1] / 1e6:.1f} MB) per step, C++ restatement of the reference CPU chain (the Rust reference cannot be built here)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "keep_rate": float(r.keep.mean())}
    emit(line)


class _OracleParquetPipeline:
    """CPU arm of config 1: the oracle behind the interface parquet_io.run_parquet drives (run_packed -> keep, content,
    metadata) -- the in-process executor of the reference without the queue (src/executor.rs:30-57)."""

    def __init__(self, steps, threads):
        import oracle

        self.oracle, self.pl, self.threads, self.decode = oracle, oracle.OraclePipeline(steps), threads, False

    def set_option(self, option, value):
        from textblaster_b200 import _abi

        if option == _abi.TB_OPT_DECODE_HTML_ENTITIES:
            self.decode = bool(value)

    def run_packed(self, values, off):
        values, off = np.asarray(values), np.asarray(off, dtype=np.int64)
        r = self.pl.run_packed(values, off, nthreads=self.threads)

        class R:
            keep = r.keep
            content = True

            def metadata_and_reason(self, i):
                return r.metadata(i), r.reason(i)

            def content_of(self, i):
                return r.content(i)

        return R()


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def run_config1(args, rank):
    """BASELINE.json configs[0]: C4QualityFilter with the parameters of the reference's own unit tests
    (c4_filters.rs:578-591 = the shipped defaults) over a 10 000-document synthetic Parquet file (length log-normal, median
    2 KB, sigma 1, clipped to [64 B, 64 KB]) through the in-process executor: file -> reader -> filter -> the two output files
    (textblaster_b200/parquet_io.run_parquet).  --impl reference: the oracle on the host cores (1 thread and all of them);
    default: the same call with the GPU pipeline.  The file round trip (pyarrow) is inside the timed region of both."""
    if rank != 0:
        return
    import shutil
    import tempfile

    import pyarrow as pa
    import pyarrow.parquet as pq

    from textblaster_b200 import parquet_io, synth
    from textblaster_b200.config import C4QualityParams

    n = 10_000
    steps = [C4QualityParams(True, True, True, 5, 3, 1000, True, True, True, True)]
    utf8, off = synth.generate(n, seed=synth.BASE_SEED + 1)
    tmp = tempfile.mkdtemp(prefix="tb_config1_")
    try:
        src = os.path.join(tmp, "in.parquet")
        col = pa.Array.from_buffers(pa.large_string(), n, [None, pa.py_buffer(off.astype(np.int64)), pa.py_buffer(utf8)])
        pq.write_table(pa.table({"id": pa.array(["doc-%d" % i for i in range(n)]), "text": col.cast(pa.string())}), src)
        cfg = parquet_io.ParquetInputConfig(src)
        threads = os.cpu_count() or 1

        def run(pipe, reps):
            best = None
            counts = None
            for k in range(reps):
                t0 = time.perf_counter()
                counts = parquet_io.run_parquet(pipe, cfg, os.path.join(tmp, "kept%d.parquet" % k), os.path.join(tmp, "exc%d.parquet" % k),
                                                decode_entities=False)
                dt = time.perf_counter() - t0
                best = dt if best is None else min(best, dt)
            return best, counts

        reps = max(1, args.steps)
        if args.impl == "reference":
            t1, c1 = run(_OracleParquetPipeline(steps, 1), max(1, min(reps, 2)))
            tn, cn = run(_OracleParquetPipeline(steps, threads), reps)
            assert c1 == cn
            pl = _OracleParquetPipeline(steps, threads)
            t0 = time.perf_counter()
            pl.pl.run_packed(utf8, off, nthreads=threads)
            t_filter = time.perf_counter() - t0
            v = n / tn
            line = {"impl": "reference", "metric": "docs/s through C4QualityFilter, Parquet file in, two Parquet files out", "value": v, "unit": UNIT,
                    "n_gpus": 0, "steps": reps, "warmup": 0, "ms_per_step": tn * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                    "dtype": "u8", "data": "synthetic",
                    "config": {"workload": "BASELINE.json configs[0]: C4QualityFilter (c4_filters.rs:578-591 parameters) on a 10k-document synthetic Parquet file via the in-process executor on the CPU",
                               "docs": n, "bytes": int(off[-1]), "cpu": cpu_model(), "cores": threads, "received_kept_excluded": list(cn),
                               "docs_per_s_1_thread": n / t1, "docs_per_s_filter_only_all_threads": n / t_filter},
                    "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                     "sample": "the whole 10k-document file, best of %d runs; C++ restatement of the reference CPU chain (the Rust reference cannot be built here)" % reps},
                    "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
            emit(line)
            return
        from textblaster_b200.executor import GpuBatchPipeline

        pipe = GpuBatchPipeline(steps, device=0)
        run(pipe, 1)  # warm-up: buffers, kernels
        tg, cg = run(pipe, reps)
        res = pipe.run_packed(utf8, off)
        t0 = time.perf_counter()
        res = pipe.run_packed(utf8, off)
        t_filter = time.perf_counter() - t0
        line = {"metric": "docs/s through C4QualityFilter, Parquet file in, two Parquet files out", "value": n / tg, "unit": UNIT, "n_gpus": 1, "steps": reps,
                "warmup": 1, "ms_per_step": tg * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": "BASELINE.json configs[0]: C4QualityFilter (c4_filters.rs:578-591 parameters) on a 10k-document synthetic Parquet file via the in-process executor, filter on the GPU",
                           "docs": n, "bytes": int(off[-1]), "received_kept_excluded": list(cg),
                           "note": "the file round trip (pyarrow read, per-document metadata, two writers) is host work inside the timed region and dominates it",
                           "docs_per_s_filter_only": n / t_filter},
                "e2e": {"value": n / tg, "unit": UNIT, "h2d_bytes_per_step": int(off[-1]) + 4 * (n + 1), "d2h_bytes_per_step": None},
                "gpu_launches": int(res.gpu_launches)}
        if not args.no_cpu_baseline:
            import oracle

            ores = oracle.OraclePipeline(steps).run_packed(utf8, off.astype(np.int64), nthreads=threads)
            ob, oo = ores.content_packed()
            mism = int((res.keep != ores.keep).sum()) + int(not (np.array_equal(res.content_offsets, oo) and np.array_equal(res.content, ob)))
            for name in res.stats[0].dtype.names:
                mism += int((res.stats[0][name] != ores.stats[0][name]).sum())
            line["parity"] = {"docs": n, "mismatches": mism, "checked": "keep, every C4 counter, rewritten content vs the oracle; tolerance 0"}
        pipe.close()
        emit(line)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def run_config3(args, rank):
    """BASELINE.json configs[2]: GopherRepetitionFilter alone on 20 000 long documents (10-200 KB, uniform in the logarithm),
    one n-gram order at a time (n = 2..10) for the top-n-gram list (fraction 0.2) and the duplicate n-gram list (0.15); per
    order docs/s and text GB/s end to end from pinned host memory, device time, and the share of documents whose tables do
    not fit shared memory (tier X, global work area).  One JSON line: `value` is the rate over all 18 sweeps together,
    `rows` the sweeps.  cpu_baseline / parity: the oracle on a bounded sample, all nine orders in both lists at once."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import torch

    import config3_sweep
    from textblaster_b200 import _abi
    from textblaster_b200.config import GopherRepetitionParams
    from textblaster_b200.executor import GpuBatchPipeline

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    utf8, off = config3_sweep.corpus()
    n, nbytes = len(off) - 1, int(off[-1])
    sampler = ClockSampler(0)
    sampler.start()
    rows = []
    total_s = 0.0
    total_gpu_ms = 0.0
    launches = 0
    ktot = {}
    for kind in ("top_n_grams", "dup_n_grams"):
        for order in range(2, 11):
            p = GpuBatchPipeline([GopherRepetitionParams(**{kind: [(order, 0.2 if kind == "top_n_grams" else 0.15)]})])
            if args.bulk_stage_tiles:
                p.set_option(_abi.TB_OPT_BULK_STAGE_TILES, args.bulk_stage_tiles)
            if args.chunk_ramp >= 0:
                p.set_option(_abi.TB_OPT_CHUNK_RAMP, args.chunk_ramp)
            p.set_option(_abi.TB_OPT_CHUNK_BYTES, args.chunk_mb << 20)
            for _ in range(max(args.warmup, 1)):
                p.run_packed(utf8, off)
            p.reset_kernel_times()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(max(args.steps, 1)):
                r = p.run_packed(utf8, off)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / max(args.steps, 1)
            kt = p.kernel_times()
            for k, v in kt.items():
                ktot[k] = ktot.get(k, 0.0) + v[0] / max(args.steps, 1)
            total_s += dt
            total_gpu_ms += r.gpu_ms
            launches += r.gpu_launches
            big = np.maximum(r.stats[0]["n_words"], r.stats[0]["n_lines"])
            rows.append({"list": kind, "n": order, "docs_per_s": n / dt, "text_gbs": nbytes / dt / 1e9, "ms": dt * 1e3, "gpu_ms": r.gpu_ms, "kept": int(r.keep.sum())})
            tier = {"w256_w512": float((big <= 512).mean()), "c1k_c2k": float(((big > 512) & (big <= 2048)).mean()),
                    "c4k_c8k": float(((big > 2048) & (big <= 8192)).mean()), "X_global": float((big > 8192).mean())}
            p.close()
    clocks = sampler.stop()
    peak, peak_src = measured_peaks()
    dom = max(ktot.items(), key=lambda kv: kv[1])
    sweeps = len(rows)
    alg = nbytes + 8 * (n + 1) + 16 * n
    line = {"metric": "docs/s through GopherRepetitionFilter, long documents", "value": n * sweeps / total_s, "unit": UNIT, "n_gpus": 1,
            "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": total_s / sweeps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "BASELINE.json configs[2]: GopherRepetitionFilter n-gram sweep (n = 2..10, top and duplicate lists, one order at a time) on %d documents of 10-200 KB (%.2f GB)" % (n, nbytes / 1e9),
                       "tier_share": tier, "spill_rate": tier["X_global"],
                       "l2": "inputs (%.2f GB per sweep) larger than the 126 MB L2" % (nbytes / 1e9)},
            "text_gbs": nbytes * sweeps / total_s / 1e9, "rows": rows,
            "e2e": {"value": n * sweeps / total_s, "unit": UNIT, "h2d_bytes_per_step": nbytes + 8 * (n + 1), "d2h_bytes_per_step": n * (2 + 248),
                    "text_gbs": nbytes * sweeps / total_s / 1e9, "ms_per_step": total_s / sweeps * 1e3},
            "gpu_launches": launches, "gpu_ms_per_step": total_gpu_ms / sweeps,
            "roofline": {"bound": "hbm", "kernel": dom[0], "achieved": alg / (dom[1] / sweeps * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg / (dom[1] / sweeps * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg, "avg_launch_ms": dom[1] / sweeps,
                         "note": "one launch per sweep processes the whole corpus; tier X runs beside the other tiers on a second stream, so the per-kernel times overlap"},
            "kernels_ms_per_step": {k: round(v / sweeps, 3) for k, v in sorted(ktot.items(), key=lambda kv: -kv[1])},
            "clocks": clocks}
    if not args.no_cpu_baseline:
        import oracle
        from parity_util import executed

        orders = list(range(2, 11))
        params = GopherRepetitionParams(dup_line_frac=0.3, dup_para_frac=0.3, dup_line_char_frac=0.2, dup_para_char_frac=0.2,
                                        top_n_grams=[(k, 0.2) for k in orders], dup_n_grams=[(k, 0.15) for k in orders])
        threads = os.cpu_count() or 1
        sample = min(n, 16 * threads)
        t0 = time.perf_counter()
        ores = oracle.OraclePipeline([params]).run_packed(utf8[:int(off[sample])], off[:sample + 1].astype(np.int64), nthreads=threads)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": sample / dt, "unit": UNIT, "cores": threads, "kind": "port", "text_gbs": int(off[sample]) / dt / 1e9,
                                "sample": "first %d documents (%.1f MB), all nine orders in both lists at once, %.1f s, oracle/tb_oracle.cpp with %d threads" % (sample, off[sample] / 1e6, dt, threads)}
        res = GpuBatchPipeline([params]).run_packed(utf8[:int(off[sample])], off[:sample + 1])
        mism = int(((res.keep != ores.keep) | (res.failed_step != ores.failed_step)).sum())
        ran = executed(ores.failed_step, 0)
        for name in res.stats[0].dtype.names:
            ne = res.stats[0][name][ran] != ores.stats[0][name][ran]
            mism += int((ne.any(axis=1) if ne.ndim > 1 else ne).sum())
        line["parity"] = {"docs": sample, "mismatches": mism, "checked": "keep, failed_step, every statistic vs the oracle on the cpu_baseline sample (the whole corpus: tools/config3_parity.py, profiles/r02_config3_parity.json)"}
    emit(line)


def run_config4(args, rank, local_rank, world):
    """BASELINE.json configs[3]: the quality chain over a stream of row groups of 65 536 documents dealt round-robin to
    the ranks, through the product's streaming runner (textblaster_b200/runner.py): several pipelines
    per GPU fed from pinned host buffers, results drained in row-group order, kept documents' content rebuilt on the host.
    The CPU generator sustains a small fraction of what one GPU consumes, so the row groups a rank streams are generated
    before the timed region (a pool of distinct groups in pinned memory; --docs-total beyond the pool walks it again: every
    pass copies the text to the device anew, nothing stays resident between groups)."""
    import torch
    import torch.distributed as dist

    from textblaster_b200 import _abi, shard, synth
    from textblaster_b200.executor import GpuBatchPipeline
    from textblaster_b200.runner import RowGroup, StreamingRunner

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    numa = pin_to_gpu_socket(torch, local_rank) if world > 1 else "single process, not pinned"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    steps = shipped_chain(args.with_fineweb, args.order)
    pool_groups = args.pool_groups
    # by default four passes over every rank's pool (each pass copies the text to the device anew): long enough for the
    # start and the end of the stream not to weigh
    total_groups = max(world, (args.docs_total or 4 * pool_groups * world * GROUP_DOCS) // GROUP_DOCS)
    mine = shard.groups_of_rank(total_groups, rank, world)  # round-robin: rank r takes groups r, r + world, ...
    threads = max(1, (os.cpu_count() or 1) // world)
    # ---- the pool: this rank's first `pool_groups` row groups, each generated from its own document indices
    t0 = time.perf_counter()
    pool = []
    for k in range(min(pool_groups, len(mine))):
        g = mine[k]
        holder = {}

        def pinned(nbytes, holder=holder):
            holder["t"] = torch.empty(max(nbytes, 1), dtype=torch.uint8, pin_memory=True)
            return holder["t"].numpy()

        u, o = synth.generate(GROUP_DOCS, seed=synth.BASE_SEED + 4, first_index=g * GROUP_DOCS, nthreads=threads, out=pinned)
        pool.append((u, o, holder))
    gen_s = time.perf_counter() - t0
    pool_bytes = sum(int(o[-1]) for _, o, _ in pool)
    gen_gbs = pool_bytes / gen_s / 1e9

    # (every pipeline gathers the kept content of its own batch: the rank's host threads are shared out among them)
    gather_threads = args.gather_threads or max(1, min(8, threads // max(args.pipelines, 1)))
    runner = StreamingRunner(lambda: GpuBatchPipeline(steps, device=local_rank), pipelines=args.pipelines,
                             copy_stats=False, gather_kept=gather_threads if any(p.CTYPE == _abi.TB_STEP_C4_QUALITY for p in steps) else 0)

    def source(n_groups):
        for k in range(n_groups):
            u, o, _ = pool[k % len(pool)]
            yield RowGroup(mine[k], u, o)

    # warm-up: the result sets and device buffers of every pipeline grow to their size, kernels are loaded
    big = max(pool, key=lambda p: int(p[1][-1]))
    runner.prime(RowGroup(-1, big[0], big[1]))
    runner.run(source(min(len(mine), max(2 * args.pipelines, 2))))
    kept_by_group = {}

    def sink(index, s):
        kept_by_group[index] = int(s.keep.sum())

    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    tot = runner.run(source(len(mine)), sink)
    barrier()
    dt = time.perf_counter() - t0
    clocks = sampler.stop()
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    agg = torch.tensor([tot.docs, tot.bytes, tot.kept, tot.kept_bytes, tot.gpu_launches], dtype=torch.int64, device="cuda")
    gms = torch.tensor([tot.gpu_ms, tot.source_wait_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(agg, op=dist.ReduceOp.SUM)
        dist.all_reduce(gms, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    docs, nbytes, kept, kept_bytes, launches = (int(x) for x in agg.tolist())
    # ---- parity on this rank's first row groups (rank 0): the same groups through a fresh pipeline with statistics, and the oracle
    parity = None
    if rank == 0 and not args.no_cpu_baseline:
        import oracle
        from parity_util import executed

        check = min(args.parity_groups, len(pool))
        pipe = GpuBatchPipeline(steps, device=local_rank)
        mism = values = ndocs = 0
        for k in range(check):
            u, o, _ = pool[k]
            res = pipe.run_packed(u, o)
            ores = oracle.OraclePipeline(steps).run_packed(u, o.astype(np.int64), nthreads=os.cpu_count() or 1)
            mism += int(((res.keep != ores.keep) | (res.failed_step != ores.failed_step)).sum())
            values += 2 * res.n_docs
            for s_i in range(len(steps)):
                ran = executed(ores.failed_step, s_i)
                for name in res.stats[s_i].dtype.names:
                    ne = res.stats[s_i][name][ran] != ores.stats[s_i][name][ran]
                    mism += int((ne.any(axis=1) if ne.ndim > 1 else ne).sum())
                    values += int(ran.sum())
            if res.content is not None:
                ob, oo = ores.content_packed()
                mism += 0 if (np.array_equal(res.content_offsets, oo) and np.array_equal(res.content, ob)) else 1
                values += res.n_docs
            mism += 0 if kept_by_group.get(mine[k]) == int(ores.keep.sum()) else 1  # what the streamed run reported for this group
            ndocs += res.n_docs
        pipe.close()
        parity = {"docs": ndocs, "values": values, "mismatches": mism,
                  "checked": "first %d row groups of rank 0: GPU vs oracle (keep, failed_step, statistics, content) and the streamed run's kept count" % check}
    runner.close()
    if rank == 0:
        rate = docs / dt
        line = {"metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": world, "steps": 1, "warmup": 1, "ms_per_step": dt * 1e3,
                "higher_is_better": True, "scaling": "weak" if not args.docs_total else "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": "full quality chain (%s order) over a stream of row groups of %d synthetic documents dealt round-robin to the ranks "
                                       "(BASELINE.json configs[3]); %d documents in all, %d row groups per rank from a pool of %d distinct pre-generated "
                                       "groups per rank (pinned host memory)" % (args.order, GROUP_DOCS, docs, len(mine), len(pool)),
                           "runner": "textblaster_b200/runner.py: %d pipelines per GPU, chunk ramps off, content as runs + kept documents gathered on the host (%d threads per pipeline)" % (args.pipelines, gather_threads),
                           "generator": "tb_synth on %d host threads sustains %.2f GB/s per rank (measured while filling the pool): far below one GPU's intake, so generation is outside the timed region" % (threads, gen_gbs),
                           "host": numa, "l2": "every row group (%.0f MB) is copied from the host again each time it is streamed" % (pool_bytes / max(len(pool), 1) / 1e6)},
                "text_gbs": nbytes / dt / 1e9, "total_docs_per_step": docs, "keep_rate": kept / max(docs, 1), "kept_bytes": kept_bytes,
                "gpu_ms_per_step": float(gms[0].item()), "source_wait_s_max": float(gms[1].item()),
                "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": nbytes + 8 * docs, "d2h_bytes_per_step": None, "text_gbs": nbytes / dt / 1e9,
                        "ms_per_step": dt * 1e3, "host_gathered_bytes_per_step": kept_bytes},
                "gpu_launches": launches, "clocks": clocks}
        if parity is not None:
            line["parity"] = parity
        emit(line)
        if parity and parity["mismatches"]:
            sys.exit(1)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--docs", type=int, default=1_000_000, help="documents per GPU per step")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--chunk-mb", type=int, default=1024)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--decode-entities", action="store_true", help="also run the HTML-entity pre-pass of the reader (SURVEY 8f row 1); not part of the headline workload")
    ap.add_argument("--order", default="baseline", choices=["baseline", "shipped"], help="step order: BASELINE.json configs[1] (C4, GopherQuality, GopherRepetition; the headline) or the shipped pipeline_config.yaml (GopherRepetition, GopherQuality, C4)")
    ap.add_argument("--content", default="auto", choices=["auto", "runs", "bytes"],
                    help="how C4's rewritten content comes back: runs of the input text (TB_OPT_RETURN_CONTENT = 2; the end-to-end step "
                         "then gathers the kept documents' content on the host, inside the timed region) or the bytes themselves (1); auto: runs, "
                         "except that the end-to-end step takes the bytes when the host has fewer than 6 cores per rank to gather with")
    ap.add_argument("--config", type=int, default=2, choices=[1, 2, 3, 4], help="BASELINE.json configuration: 1 (C4 on a 10k-document Parquet file), 2 (headline, default), 3 (n-gram sweep on long documents), 4 (streamed row groups)")
    ap.add_argument("--docs-total", type=int, default=0, help="config 4: documents over all ranks (0: four passes over every rank's pool)")
    ap.add_argument("--pool-groups", type=int, default=16, help="config 4: distinct row groups generated per rank before the timed region")
    ap.add_argument("--pipelines", type=int, default=4, help="config 4: pipelines (and host threads) per GPU")
    ap.add_argument("--parity-groups", type=int, default=2, help="config 4: row groups of rank 0 checked against the oracle")
    ap.add_argument("--host-memory", default="pinned", choices=["pinned", "pageable"],
                    help="where the text of the end-to-end step lives: pinned host memory (default) or pageable memory as Arrow allocates it "
                         "(the library then stages it through its own pinned ring, TB_OPT_STAGE_THREADS)")
    ap.add_argument("--stage-threads", type=int, default=-1, help="TB_OPT_STAGE_THREADS for --host-memory pageable (-1: library default, 0: off)")
    ap.add_argument("--bulk-stage-tiles", type=int, default=0, help="TB_OPT_BULK_STAGE_TILES: bulk-copy staging of documents of at least this many 1 KiB tiles (0: off)")
    ap.add_argument("--e2e-api", default="best", choices=["best", "runner", "call"],
                    help="how the end-to-end figure of the headline is taken: one tb_pipeline_run_batch call per step (call), the step's documents "
                         "as equal row groups streamed through StreamingRunner with --pipelines pipelines per GPU (runner), or both and the faster "
                         "of the two as the figure, the other beside it (best, the default: the stream wins when the chain starts with the cheap C4 "
                         "passes, the single call when it starts with the repetition filter)")
    ap.add_argument("--stream-group-docs", type=int, default=0, help="headline, streamed end-to-end figure: documents per row group (0: about %d)" % GROUP_DOCS)
    ap.add_argument("--chunk-ramp", type=int, default=-1, help="TB_OPT_CHUNK_RAMP (-1: library default)")
    ap.add_argument("--gather-threads", type=int, default=0, help="host threads of TB_OPT_GATHER_KEPT (0: min(8, cpus / ranks))")
    ap.add_argument("--with-fineweb", action="store_true",
                    help="append FineWebQualityFilter (SURVEY 8(f) row 3) to the chain; not the headline workload")
    args = ap.parse_args()
    guard_stdout()
    args.warmup = max(args.warmup, 0)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.config == 1:
        run_config1(args, rank)
        return
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.config == 3:
        run_config3(args, rank)
        return
    if args.config == 4:
        if args.order == "baseline" and "--order" not in sys.argv:
            args.order = "shipped"  # config 4 is the shipped pipeline_config.yaml chain unless asked otherwise
        run_config4(args, rank, local_rank, world)
        return

    import torch
    import torch.distributed as dist

    from textblaster_b200 import _abi, synth
    from textblaster_b200.executor import GpuBatchPipeline

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    numa = pin_to_gpu_socket(torch, local_rank) if world > 1 else "single process, not pinned"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- this rank's row groups of the synthetic corpus (round-robin, no overlap between ranks), in pinned host memory
    from textblaster_b200 import shard

    t_gen = time.perf_counter()
    groups = shard.row_groups(world * args.docs, min(GROUP_DOCS, args.docs))
    mine = shard.groups_of_rank(len(groups), rank, world)
    parts = [synth.generate(b - a, seed=synth.BASE_SEED + 2, first_index=a) for a, b in (groups[g] for g in mine)]
    n_docs = sum(len(o) - 1 for _, o in parts)
    n_bytes = sum(int(o[-1]) for _, o in parts)
    holder = {"t": torch.empty(max(n_bytes, 1), dtype=torch.uint8, pin_memory=True)}
    utf8 = holder["t"].numpy()
    off_t = torch.empty(n_docs + 1, dtype=torch.int64, pin_memory=True)
    off = off_t.numpy()
    pos = doc = 0
    for u, o in parts:
        utf8[pos:pos + int(o[-1])] = u[:int(o[-1])]
        off[doc:doc + len(o)] = o + pos
        pos += int(o[-1]); doc += len(o) - 1
    del parts
    t_gen = time.perf_counter() - t_gen
    steps = shipped_chain(args.with_fineweb, args.order)
    pipe = GpuBatchPipeline(steps, device=local_rank)
    pipe.set_option(_abi.TB_OPT_CHUNK_BYTES, args.chunk_mb << 20)
    if args.decode_entities:
        pipe.set_option(_abi.TB_OPT_DECODE_HTML_ENTITIES, 1)
    if args.bulk_stage_tiles:
        pipe.set_option(_abi.TB_OPT_BULK_STAGE_TILES, args.bulk_stage_tiles)
    runs_mode = args.content in ("runs", "auto") and not args.decode_entities
    # the end-to-end step rebuilds the kept documents' content on host threads: with few cores per rank (many GPUs on
    # one host) the copy engine is the cheaper way to get the content into host memory, so `auto` takes the bytes there
    e2e_runs_mode = runs_mode and (args.content == "runs" or (os.cpu_count() or 1) // world >= 6)
    pipe.set_option(_abi.TB_OPT_RETURN_CONTENT, 2 if runs_mode else 1)
    # runs mode: the end-to-end call also rebuilds the content of the kept documents (what the writer of the output
    # file takes) on host threads inside the library, chunk by chunk behind the device
    gather_threads = args.gather_threads or max(1, min(8, (os.cpu_count() or 1) // world))
    if runs_mode:
        pipe.set_option(_abi.TB_OPT_GATHER_KEPT, gather_threads)
    d_text = holder["t"].cuda(non_blocking=False)
    d_off = off_t.cuda()
    out = _abi.tb_batch_out()

    def content_bytes_back():
        return int(out.n_runs) * 8 if out.content_kind == 2 else int(out.content_bytes)

    def step_device():
        pipe.run_device_raw(d_text.data_ptr(), d_off.data_ptr(), n_docs, n_bytes, out)
        r = (out.gpu_launches, out.gpu_ms, int(np.ctypeslib.as_array(out.keep, (n_docs,)).sum()) if n_docs else 0, content_bytes_back())
        pipe.release(out)
        return r

    last_keep = {}

    host_utf8 = utf8
    if args.host_memory == "pageable":
        host_utf8 = np.array(utf8, copy=True)  # ordinary (pageable) memory, as pyarrow hands a column over
        if args.stage_threads >= 0:
            pipe.set_option(_abi.TB_OPT_STAGE_THREADS, args.stage_threads)

    def step_host():
        pipe.run_raw(host_utf8.ctypes.data, off.ctypes.data, True, n_docs, out)
        keep = np.ctypeslib.as_array(out.keep, (n_docs,)) if n_docs else np.zeros(0, np.uint8)
        last_keep["k"] = keep.copy()
        if out.content_kind == 2:  # the content of the documents that are written out was rebuilt from the input column
            nk = int(out.n_kept)
            assert nk == int(keep.sum()) and out.kept_offsets[0] == 0
            last_keep["kept_bytes"] = int(out.kept_offsets[nk])
        r = (out.gpu_launches, out.gpu_ms, int(keep.sum()), content_bytes_back())
        pipe.release(out)
        return r

    def timed(fn):
        for _ in range(args.warmup):
            fn()
        pipe.reset_kernel_times()
        barrier()
        sampler = ClockSampler(local_rank)
        if rank == 0:  # one poller for the job: concurrent nvidia-smi queries on every rank disturb the launches
            sampler.start()
        launches = 0
        gpu_ms = 0.0
        t0 = time.perf_counter()
        for _ in range(args.steps):
            l, ms, kept, cbytes = fn()
            launches += l
            gpu_ms += ms
        barrier()
        dt = time.perf_counter() - t0
        clocks = sampler.stop()
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), launches, gpu_ms, kept, cbytes, clocks, pipe.kernel_times()

    def copy_rate(back):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        if back:
            holder["t"].copy_(d_text, non_blocking=True)
        else:
            d_text.copy_(holder["t"], non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        return holder["t"].numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9

    def h2d_probe(back=False):
        """Pinned host -> device copy rate of this box (the step's own text buffer, CUDA events, best of 3);
        back=True: device -> pinned host, the way the results leave."""
        return max(copy_rate(back) for _ in range(3))

    def concurrent_probe(back=False):
        """The same copy on every rank at once (a barrier before each of 3 rounds): the sum of the ranks' rates,
        median over the rounds = what the host can move for all GPUs of the job together."""
        sums = []
        for _ in range(3):
            barrier()
            t = torch.tensor([copy_rate(back)], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            sums.append(float(t.item()))
        return sorted(sums)[1]

    pinned = holder["t"].is_pinned()
    h2d_peak = h2d_probe() if pinned else None
    d2h_peak = h2d_probe(back=True) if pinned else None
    h2d_all = concurrent_probe() if world > 1 and pinned else None
    d2h_all = concurrent_probe(back=True) if world > 1 and pinned else None
    dt_dev, launches, gpu_ms, kept, cbytes, clocks, ktimes = timed(step_device)
    if runs_mode and not e2e_runs_mode:
        pipe.set_option(_abi.TB_OPT_RETURN_CONTENT, 1)
        pipe.set_option(_abi.TB_OPT_GATHER_KEPT, 0)
    dt_e2e, launches_e, gpu_ms_e, _, cbytes_e, clocks_e, _ = timed(step_host)
    # ---- the same steps as a stream of row groups through the runner (textblaster_b200/runner.py): several pipelines of
    # one context take the rank's row groups in turn, so the copies of one batch and the kernel tails of another overlap
    single_call = {"value": None, "ms_per_step": dt_e2e / max(args.steps, 1) * 1e3, "d2h_content_bytes_per_step": cbytes_e}
    use_runner = (args.e2e_api in ("runner", "best") and args.host_memory == "pinned" and not args.decode_entities and not args.bulk_stage_tiles
                  and n_docs >= 4 * min(GROUP_DOCS, args.docs))
    stream = None
    if use_runner:
        from textblaster_b200.runner import RowGroup, StreamingRunner
        P = max(1, args.pipelines)
        gt = (args.gather_threads or max(1, min(8, ((os.cpu_count() or 1) // world) // P))) if e2e_runs_mode else 0
        # the rank's documents as equal row groups of about GROUP_DOCS (a short last group would have the pipelines fall
        # out of step: the results are drained in order)
        G = max(1, round(n_docs / min(args.stream_group_docs or GROUP_DOCS, args.docs)))
        bounds = [(n_docs * k // G, n_docs * (k + 1) // G) for k in range(G)]

        def make():
            return GpuBatchPipeline(steps, device=local_rank)  # (content as bytes unless the runner is told to gather: gather_kept)

        runner = StreamingRunner(make, pipelines=P, copy_stats=False, gather_kept=gt)

        copies = None
        if os.environ.get("TB_BENCH_STREAM_COPY") == "1":  # (diagnosis: every row group in a pinned buffer of its own, offsets from 0)
            copies = []
            for a, b in bounds:
                tu = torch.empty(int(off[b] - off[a]) + 64, dtype=torch.uint8, pin_memory=True)
                to = torch.empty(b - a + 1, dtype=torch.int64, pin_memory=True)
                tu.numpy()[:int(off[b] - off[a])] = utf8[int(off[a]):int(off[b])]
                to.numpy()[:] = off[a:b + 1] - off[a]
                copies.append((tu, to))

        def source(n_steps):
            for s_ in range(n_steps):
                for k, (a, b) in enumerate(bounds):
                    if copies:
                        yield RowGroup(s_ * len(bounds) + k, copies[k][0].numpy()[:int(off[b] - off[a])], copies[k][1].numpy())
                    else:
                        yield RowGroup(s_ * len(bounds) + k, utf8, off[a:b + 1])

        big = max(bounds, key=lambda ab: int(off[ab[1]] - off[ab[0]]))
        runner.prime(RowGroup(-1, utf8, off[big[0]:big[1] + 1]))
        stream_dbg = {}
        runner.run(source(max(args.warmup, 1)))
        stream_keep = np.zeros(n_docs, np.uint8)
        stream_tot = {"kept_bytes": 0, "runs_bytes": 0}

        def sink(index, sm):
            a, b = bounds[index % len(bounds)]
            if index < len(bounds):
                stream_keep[a:b] = sm.keep
            if sm.kept_offsets is not None:
                stream_tot["kept_bytes"] += int(sm.kept_offsets[-1])

        barrier()
        sampler_s = ClockSampler(local_rank)
        if rank == 0:
            sampler_s.start()
        t0 = time.perf_counter()
        tot_s = runner.run(source(args.steps), sink)
        barrier()
        dt_s = time.perf_counter() - t0
        clocks_s = sampler_s.stop()
        ts = torch.tensor([dt_s], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        dt_s = float(ts.item())
        assert np.array_equal(stream_keep, last_keep["k"]), "the streamed run and the single call disagree on the keep mask"
        stream_dbg.update(gpu_ms=tot_s.gpu_ms, source_wait_s=tot_s.source_wait_s)
        if rank == 0 and os.environ.get("TB_BENCH_DEBUG"):
            print("[bench] stream: %.1f ms/step, gpu_ms %.1f per step, source wait %.3f s" % (dt_s / args.steps * 1e3, tot_s.gpu_ms / args.steps, tot_s.source_wait_s), file=sys.stderr)
        stream = {"dt": dt_s, "groups": len(bounds), "pipelines": P, "gather_threads": gt, "launches": tot_s.gpu_launches, "clocks": clocks_s,
                  "kept_bytes_per_step": stream_tot["kept_bytes"] // max(args.steps, 1)}
        runner.close()
    # host-side gather of the keep masks onto rank 0, ordered by row group (the only cross-rank traffic of the job)
    t0 = time.perf_counter()
    local, pos = {}, 0
    for g in mine:
        a, b = groups[g]
        local[g] = last_keep["k"][pos:pos + (b - a)]
        pos += b - a
    masks = shard.gather_keep_masks(local, len(groups), rank, world)
    gather_s = time.perf_counter() - t0

    tot = torch.tensor([n_docs, n_bytes], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    total_docs, total_bytes = int(tot[0].item()), int(tot[1].item())
    if rank == 0:
        K = max(args.steps, 1)
        value = total_docs * K / dt_dev
        single_call["value"] = total_docs * K / dt_e2e
        streamed = None
        if stream:
            streamed = {"value": total_docs * K / stream["dt"], "ms_per_step": stream["dt"] / K * 1e3}
            if args.e2e_api == "best" and stream["dt"] > dt_e2e:
                stream = None  # the single call was the faster one: it is the figure, the stream stays beside it
            else:
                dt_e2e = stream["dt"]
        e2e = total_docs * K / dt_e2e
        # roofline of the dominant kernel (device-resident run)
        dom = max(ktimes.items(), key=lambda kv: kv[1][0]) if ktimes else ("", (0.0, 0))
        dom_name, (dom_ms, dom_launches) = dom
        alg_bytes_step = n_bytes + 8 * (n_docs + 1) + 16 * n_docs  # text once + i64 offsets + minimal per-doc output
        launches_per_step = max(dom_launches / K, 1)
        alg_bytes_launch = alg_bytes_step / launches_per_step
        units_note = "every document of the chunk"
        # (the repetition pass's tiers by size: tb_rep_cap in csrc/tb_kernels.cuh)
        tier_of = {"k_gr_rep_w256": 0, "k_gr_rep_w512": 1, "k_gr_rep_c1k": 2, "k_gr_rep_c2k": 3, "k_gr_rep_c4k": 4, "k_gr_rep_c8k": 5}
        if dom_name in tier_of:
            # a tier kernel of the repetition pass only works on the documents of its size class that reached the step:
            # it is credited with their text (and their 24 bytes of offsets / results), not with the whole chunk's
            gr_idx = [i for i, prm in enumerate(steps) if prm.CTYPE == _abi.TB_STEP_GOPHER_REPETITION][0]
            res = pipe.run_packed(utf8, off, copy_content=False)
            ran = (res.failed_step < 0) | (res.failed_step >= gr_idx)
            big = np.maximum(res.stats[gr_idx]["n_words"], res.stats[gr_idx]["n_lines"])
            lo, hi = [(0, 256), (256, 512), (512, 1024), (1024, 2048), (2048, 4096), (4096, 8192)][tier_of[dom_name]]
            sel = ran & (big > lo if lo else big >= 0) & (big <= hi)
            tier_bytes = int(np.diff(off)[sel].sum()) + 24 * int(sel.sum())
            alg_bytes_launch = tier_bytes / launches_per_step
            units_note = "the %d documents with %d < max(words, line pieces) <= %d that reached the step (%.2f GB of the step's text)" % (int(sel.sum()), lo, hi, tier_bytes / 1e9)
        avg_ms = dom_ms / max(dom_launches, 1)
        achieved = alg_bytes_launch / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        peak, peak_src = measured_peaks()
        stats_bytes = sum(C.sizeof(_abi.STATS_STRUCT[p.CTYPE]) for p in steps)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt_dev / K * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD + (" + FineWebQualityFilter" if args.with_fineweb else "") + (" [shipped step order]" if args.order == "shipped" else "") + (" + HTML entity pre-pass" if args.decode_entities else ""), "docs_per_gpu_per_step": n_docs, "bytes_per_gpu_per_step": n_bytes,
                       "sharding": "row groups of %d docs round-robin over ranks, no data-path collective; keep masks gathered on the host (%.3f s, outside the timed region)" % (min(GROUP_DOCS, args.docs), gather_s),
                       "kept_total": int(sum(int(m.sum()) for m in masks)),
                       "content": ("runs of the input text (TB_OPT_RETURN_CONTENT = 2); the e2e call also rebuilds the content of the kept documents "
                                   "on the host from the runs and the input column (TB_OPT_GATHER_KEPT, %d threads, %.2f GB per step), "
                                   "inside the timed region" % (gather_threads, last_keep.get("kept_bytes", 0) / 1e9))
                                  if e2e_runs_mode else
                                  ("value: runs of the input text (TB_OPT_RETURN_CONTENT = 2); e2e: the bytes (TB_OPT_RETURN_CONTENT = 1) -- %d host cores per rank "
                                   "are too few to rebuild the kept documents' content behind the device" % ((os.cpu_count() or 1) // world))
                                  if runs_mode else "bytes (TB_OPT_RETURN_CONTENT = 1)",
                       "host_memory": args.host_memory + ("" if args.host_memory == "pinned" else " (staged by the library, TB_OPT_STAGE_THREADS = %s)" % ("default 4" if args.stage_threads < 0 else args.stage_threads)),
                       "chunk_mb": args.chunk_mb, "host": numa, "host_cpus": os.cpu_count(), "l2": "inputs (%.2f GB per step) larger than the 126 MB L2" % (n_bytes / 1e9),
                       "generator_s": round(t_gen, 2)},
            "text_gbs": total_bytes * K / dt_dev / 1e9, "total_docs_per_step": total_docs,
            "gpu_ms_per_step": gpu_ms / K, "keep_rate": kept / max(n_docs, 1),
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": n_bytes + 8 * (n_docs + 1),
                    "d2h_bytes_per_step": n_docs * (2 + stats_bytes) + cbytes_e + 8 * (n_docs + 1), "text_gbs": total_bytes * K / dt_e2e / 1e9,
                    "ms_per_step": dt_e2e / K * 1e3,
                    "content": "runs + host gather" if e2e_runs_mode else "bytes",
                    "host_gathered_bytes_per_step": last_keep.get("kept_bytes") if e2e_runs_mode else None,
                    "api": ("StreamingRunner.run (textblaster_b200/runner.py): the step's documents as %d equal row groups streamed through %d pipelines of "
                            "one context, %d gather threads each; results drained in row-group order; keep mask equal to the single call's; the "
                            "stream overlaps the copies and the kernel tails of %d batches, which is why it is faster than `value` (one batch at a "
                            "time on one pipeline, where the per-kernel times and the roofline are taken)"
                            % (stream["groups"], stream["pipelines"], stream["gather_threads"], stream["pipelines"])) if stream
                           else "tb_pipeline_run_batch: one call per step",
                    "single_call": single_call if stream else None, "streamed": streamed if not stream else None},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": dom_name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak if peak else None, "traffic": ncu_traffic(dom_name, alg_bytes_step / launches_per_step), "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg_bytes_launch, "avg_launch_ms": avg_ms, "units": units_note,
                         "share_of_gpu_time": dom_ms / max(sum(v[0] for v in ktimes.values()), 1e-9)},
            # the whole chain against HBM (algorithmic bytes / sum of kernel time) and the end-to-end
            # run against the host-to-device link (bytes sent / wall time of the copy-overlapped pipeline)
            "chain_hbm": {"achieved": alg_bytes_step * K / max(gpu_ms * 1e-3, 1e-12) / 1e9, "peak": peak, "unit": "GB/s",
                          "frac": (alg_bytes_step * K / max(gpu_ms * 1e-3, 1e-12) / 1e9) / peak if peak else None},
            "h2d_roofline": {"achieved": (n_bytes + 8 * (n_docs + 1)) * K / dt_e2e / 1e9, "peak": h2d_peak, "unit": "GB/s",
                             "frac": ((n_bytes + 8 * (n_docs + 1)) * K / dt_e2e / 1e9) / h2d_peak if h2d_peak else None,
                             "peak_source": "pinned host->device copy of this step's text buffer, measured in this run (best of 3)",
                             "per_gpu": True,
                             # whole job: bytes all ranks sent per second against the sum of the per-rank rates of the
                             # probe run on all ranks at once (N > 1)
                             "job_achieved": (total_bytes + 8 * (total_docs + world)) * K / dt_e2e / 1e9,
                             "job_peak_concurrent": h2d_all,
                             "job_frac": ((total_bytes + 8 * (total_docs + world)) * K / dt_e2e / 1e9) / h2d_all if h2d_all else None},
            # the results (keep, statistics, content) go back over the same link in the other direction; on a box whose
            # host cannot take the results of all its GPUs at once this, not the kernels, bounds the job (N = 8)
            "d2h_roofline": {"achieved": (n_docs * (2 + stats_bytes) + cbytes + 8 * (n_docs + 1)) * K / dt_dev / 1e9, "peak": d2h_peak, "unit": "GB/s",
                             "frac": ((n_docs * (2 + stats_bytes) + cbytes + 8 * (n_docs + 1)) * K / dt_dev / 1e9) / d2h_peak if d2h_peak else None,
                             "per_gpu": True, "measured_on": "the device-resident run (value)",
                             "job_peak_concurrent": d2h_all,
                             "job_frac_per_gpu": ((n_docs * (2 + stats_bytes) + cbytes + 8 * (n_docs + 1)) * K / dt_dev / 1e9) / (d2h_all / world) if d2h_all else None},
            "kernels_ms_per_step": {k: round(v[0] / K, 3) for k, v in sorted(ktimes.items(), key=lambda kv: -kv[1][0])},
            "clocks": clocks,
        }
        parity_failed = False
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"], ores, sample = cpu_baseline(utf8, off, steps)
            if not args.decode_entities:  # (the oracle arm sees the text as generated; the entity pre-pass has its own tests)
                line["parity"] = parity_check(pipe, steps, utf8, off, sample, ores)
                parity_failed = line["parity"]["mismatches"] != 0
        emit(line)
        if parity_failed:
            sys.stderr.write("bench.py: the GPU results differ from the oracle on the checked sample: %s\n" % json.dumps(line["parity"]))
            sys.exit(1)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
