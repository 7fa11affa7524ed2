"""Streaming runner: row groups in, outcomes out, one GPU kept busy by a stream of batches.

The reference moves one document per AMQP message from a producer that reads Parquet batches
(src/producer_logic.rs:30-66) to competing workers (src/worker_logic.rs:241-283) and back to an aggregator
(src/producer_logic.rs:109-196).  Here a rank (one process per GPU) takes its row groups from a source, keeps
``pipelines`` GpuBatchPipeline instances of one device context busy from as many host threads -- while one batch's
kernels run, the next batch's text crosses the host-to-device link -- and hands the results to a sink in row-group
order.  No collective: row groups are independent; the caller deals them to ranks
(``shard.groups_of_rank``) and gathers what it needs on the host.

The source yields ``RowGroup(index, utf8, offsets)`` with the text in host memory (pinned for full link speed).  The
sink is called as ``sink(index, summary)`` in source order; ``summary`` holds the keep mask, the failing step, the
per-step statistics and, with ``gather_kept``, the content of the kept documents.  The arrays of a summary are views
of the library's result buffers: they are valid during the sink call only (the batch is released when it returns);
a pipeline takes its next row group meanwhile with a result set of its own.
"""
from __future__ import annotations

import ctypes as C
import queue
import threading
import time
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterable, List, Optional, Sequence

import numpy as np

from . import _abi


@dataclass
class RowGroup:
    index: int
    utf8: np.ndarray      # uint8, the Arrow values buffer of the text column (host memory)
    offsets: np.ndarray   # int32 / int64, n + 1 entries
    release: Optional[Callable[[], None]] = None   # called when the buffers may be reused


@dataclass
class GroupSummary:
    index: int
    n_docs: int
    n_bytes: int
    keep: np.ndarray
    failed_step: np.ndarray
    stats: Optional[list] = None
    kept_content: Optional[np.ndarray] = None
    kept_offsets: Optional[np.ndarray] = None
    gpu_ms: float = 0.0
    gpu_launches: int = 0
    _release: Optional[Callable[[], None]] = None   # gives the result buffers back to the pipeline


@dataclass
class RunTotals:
    groups: int = 0
    docs: int = 0
    bytes: int = 0
    kept: int = 0
    kept_bytes: int = 0
    gpu_ms: float = 0.0
    gpu_launches: int = 0
    wall_s: float = 0.0
    source_wait_s: float = 0.0   # time the workers spent waiting for the source (a slow reader shows up here)
    per_step_failed: Dict[int, int] = field(default_factory=dict)


class StreamingRunner:
    def __init__(self, make_pipeline: Callable[[], object], pipelines: int = 2, copy_stats: bool = False, gather_kept: int = 0,
                 queue_depth: Optional[int] = None, results_per_pipeline: int = 2):
        """make_pipeline: () -> GpuBatchPipeline (called ``pipelines`` times; all on one device).  gather_kept: host
        threads per pipeline that rebuild the kept documents' content (TB_OPT_GATHER_KEPT; needs a C4 step).
        results_per_pipeline: result sets a pipeline may have outstanding (handed to the sink, not yet released); a
        pipeline waits for the sink beyond that, so the pinned result buffers stay a fixed set (``prime`` creates them)."""
        self.pipes = [make_pipeline() for _ in range(max(1, pipelines))]
        self.copy_stats = copy_stats
        self.gather_kept = gather_kept
        self.queue_depth = queue_depth or (len(self.pipes) + 1)
        self.results_per_pipeline = max(1, results_per_pipeline)
        self._slots = {id(p): threading.Semaphore(self.results_per_pipeline) for p in self.pipes}
        for p in self.pipes:
            if hasattr(p, "set_option"):
                p.set_option(_abi.TB_OPT_CHUNK_RAMP, 0)
                if gather_kept:
                    p.set_option(_abi.TB_OPT_RETURN_CONTENT, 2)
                    p.set_option(_abi.TB_OPT_GATHER_KEPT, gather_kept)

    def prime(self, g: RowGroup):
        """Run one (full-size) row group through every pipeline ``results_per_pipeline`` times without releasing in
        between: the library then owns that many result sets per pipeline, grown to size, before anything is timed."""
        for p in self.pipes:
            if not hasattr(p, "run_raw"):
                continue
            held = []
            n = len(g.offsets) - 1
            for _ in range(self.results_per_pipeline):
                out = _abi.tb_batch_out()
                p.run_raw(g.utf8.ctypes.data if g.utf8.size else None, g.offsets.ctypes.data, g.offsets.dtype == np.int64, n, out)
                held.append(out)
            for out in held:
                p.release(out)

    def close(self):
        for p in self.pipes:
            if hasattr(p, "close"):
                p.close()

    def _summarise(self, pipe, g: RowGroup) -> GroupSummary:
        n = len(g.offsets) - 1
        nb = int(g.offsets[-1] - g.offsets[0]) if n else 0
        if not hasattr(pipe, "run_raw"):  # a stand-in with run_packed only (CPU tests)
            r = pipe.run_packed(g.utf8, g.offsets)
            return GroupSummary(g.index, n, nb, np.asarray(r.keep).copy(), np.asarray(r.failed_step).copy(),
                                list(r.stats) if self.copy_stats else None)
        slot = self._slots[id(pipe)]
        slot.acquire()  # (released with the result set, after the sink)
        out = _abi.tb_batch_out()
        try:
            pipe.run_raw(g.utf8.ctypes.data if g.utf8.size else None, g.offsets.ctypes.data, g.offsets.dtype == np.int64, n, out)
        except BaseException:
            slot.release()
            raise
        try:
            s = GroupSummary(g.index, n, nb,
                             np.ctypeslib.as_array(out.keep, (n,)) if n else np.zeros(0, np.uint8),
                             np.ctypeslib.as_array(out.failed_step, (n,)) if n else np.zeros(0, np.int8),
                             gpu_ms=out.gpu_ms, gpu_launches=out.gpu_launches)
            if self.copy_stats:
                s.stats = []
                for k, prm in enumerate(pipe.params):
                    dt = _abi.stats_dtype(prm.CTYPE)
                    raw = (C.c_uint8 * (n * dt.itemsize)).from_address(out.step_stats[k]) if n else b""
                    s.stats.append(np.frombuffer(raw, dtype=dt) if n else np.zeros(0, dt))
            if out.kept_offsets:
                nk = int(out.n_kept)
                s.kept_offsets = np.ctypeslib.as_array(out.kept_offsets, (nk + 1,))
                kb = int(s.kept_offsets[-1])
                s.kept_content = np.ctypeslib.as_array(out.kept_content, (kb,)) if kb else np.zeros(0, np.uint8)
            s._release = lambda: (pipe.release(out), slot.release())
            return s
        except BaseException:
            pipe.release(out)
            slot.release()
            raise

    def run(self, source: Iterable[RowGroup], sink: Optional[Callable[[int, GroupSummary], None]] = None) -> RunTotals:
        tot = RunTotals()
        q: "queue.Queue[Optional[RowGroup]]" = queue.Queue(maxsize=self.queue_depth)
        done: Dict[int, GroupSummary] = {}
        order: List[int] = []
        lock = threading.Lock()
        cond = threading.Condition(lock)
        errors: List[BaseException] = []
        wait_s = [0.0]

        def put(item):  # gives up once the run has failed (the workers are gone then)
            while not errors:
                try:
                    q.put(item, timeout=0.05)
                    return True
                except queue.Full:
                    pass
            return False

        def feeder():
            try:
                for g in source:
                    with lock:
                        order.append(g.index)
                    if not put(g):
                        return
            except BaseException as e:  # a failing source stops the run
                errors.append(e)
            finally:
                for _ in self.pipes:
                    if not put(None):
                        break

        def worker(pipe):
            while True:
                t0 = time.perf_counter()
                g = None
                while not errors:
                    try:
                        g = q.get(timeout=0.05)
                        break
                    except queue.Empty:
                        pass
                dt = time.perf_counter() - t0
                if g is None:
                    return
                try:
                    s = self._summarise(pipe, g)
                except BaseException as e:
                    errors.append(e)
                    s = None
                finally:
                    if g.release:
                        g.release()
                with cond:
                    wait_s[0] += dt
                    if s is not None:
                        done[g.index] = s
                    cond.notify_all()
                if s is None:
                    return

        t_start = time.perf_counter()
        threads = [threading.Thread(target=feeder, daemon=True)] + [threading.Thread(target=worker, args=(p,), daemon=True) for p in self.pipes]
        for t in threads:
            t.start()
        # drain in source order
        emitted = 0
        while True:
            with cond:
                while not errors and not (emitted < len(order) and order[emitted] in done) and any(t.is_alive() for t in threads[1:]):
                    cond.wait(timeout=0.05)
                if errors:
                    break
                if emitted < len(order) and order[emitted] in done:
                    s = done.pop(order[emitted])
                    emitted += 1
                else:
                    if not any(t.is_alive() for t in threads[1:]) and not (emitted < len(order) and order[emitted] in done):
                        break
                    continue
            tot.groups += 1
            tot.docs += s.n_docs
            tot.bytes += s.n_bytes
            tot.kept += int(s.keep.sum())
            tot.gpu_ms += s.gpu_ms
            tot.gpu_launches += s.gpu_launches
            if s.kept_offsets is not None:
                tot.kept_bytes += int(s.kept_offsets[-1])
            fs = s.failed_step[s.failed_step >= 0]
            if fs.size:
                for k, c in zip(*np.unique(fs, return_counts=True)):
                    tot.per_step_failed[int(k)] = tot.per_step_failed.get(int(k), 0) + int(c)
            try:
                if sink:
                    sink(s.index, s)
            finally:
                if s._release:
                    s._release()
                    s._release = None
        for t in threads:
            t.join(timeout=60)
        for s in done.values():  # (a failed run leaves summaries nobody consumed)
            if s._release:
                s._release()
        if errors:
            raise errors[0]
        tot.wall_s = time.perf_counter() - t_start
        tot.source_wait_s = wait_s[0]
        return tot
