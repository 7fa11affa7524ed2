"""The N > 1 path on CPU: two gloo ranks each filter their row groups (with the CPU oracle standing in
for the GPU chain, which needs a device) and rank 0 gathers the keep masks; the result must equal a
single-process run over the whole corpus, in row order."""
import os
import socket

import numpy as np
import pytest

import oracle
from parity_cases import CHAINS
from textblaster_b200 import shard, synth

N_DOCS, GROUP = 1200, 250


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    groups = shard.row_groups(N_DOCS, GROUP)
    utf8, off = synth.generate(N_DOCS, seed=5, median_bytes=600, nthreads=1)
    sizes = [int(off[b] - off[a]) for a, b in groups]
    mine = shard.groups_of_rank(len(groups), rank, world, sizes)
    # the rank's row groups go through the product's streaming runner (two pipelines, results in row-group order);
    # the oracle stands in for the GPU pipeline, which needs a device
    from textblaster_b200.runner import RowGroup, StreamingRunner

    runner = StreamingRunner(lambda: oracle.OraclePipeline(CHAINS["baseline_order"]), pipelines=2)
    local, seen = {}, []

    def sink(index, s):
        local[index] = s.keep
        seen.append(index)

    tot = runner.run((RowGroup(g, utf8[off[groups[g][0]]:off[groups[g][1]]], off[groups[g][0]:groups[g][1] + 1] - off[groups[g][0]]) for g in mine), sink)
    assert seen == list(mine) and tot.groups == len(mine) and tot.docs == sum(groups[g][1] - groups[g][0] for g in mine)
    masks = shard.gather_keep_masks(local, len(groups), rank, world)
    if rank == 0:
        q.put((np.concatenate(masks), mine))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_matches_single_process():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got, mine0 = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    utf8, off = synth.generate(N_DOCS, seed=5, median_bytes=600, nthreads=1)
    want = oracle.OraclePipeline(CHAINS["baseline_order"]).run_packed(utf8, off).keep
    assert np.array_equal(got, want)
    assert 0 < len(mine0) < len(shard.row_groups(N_DOCS, GROUP))


def test_assignment_is_a_partition():
    for world in (1, 2, 4, 8):
        for sizes in (None, [5, 1, 9, 3, 3, 7, 2, 8, 4, 6, 1]):
            n = 11
            parts = [shard.groups_of_rank(n, r, world, sizes) for r in range(world)]
            assert sorted(g for p in parts for g in p) == list(range(n))
    assert shard.row_groups(10, 4) == [(0, 4), (4, 8), (8, 10)]


def test_streaming_runner_orders_results_and_reports_errors():
    """several pipelines finish out of order; the sink still sees the row groups in source order, totals add up,
    and a failing pipeline call surfaces as the run's exception"""
    import random
    import time

    from textblaster_b200.runner import RowGroup, StreamingRunner

    utf8, off = synth.generate(900, seed=11, median_bytes=500, nthreads=1)
    groups = shard.row_groups(900, 100)
    want = oracle.OraclePipeline(CHAINS["shipped_order"]).run_packed(utf8, off)

    class Slow:
        def __init__(self, seed):
            self.rng, self.pl = random.Random(seed), oracle.OraclePipeline(CHAINS["shipped_order"])

        def run_packed(self, u, o):
            time.sleep(self.rng.random() * 0.02)
            return self.pl.run_packed(u, o)

    seeds = iter(range(10))
    runner = StreamingRunner(lambda: Slow(next(seeds)), pipelines=3, copy_stats=True)
    got = {}
    order = []
    tot = runner.run((RowGroup(i, utf8[off[a]:off[b]], off[a:b + 1] - off[a]) for i, (a, b) in enumerate(groups)),
                     lambda i, s: (order.append(i), got.__setitem__(i, s)))
    assert order == list(range(len(groups))) and tot.docs == 900 and tot.kept == int(want.keep.sum())
    assert np.array_equal(np.concatenate([got[i].keep for i in order]), want.keep)
    assert np.array_equal(np.concatenate([got[i].stats[0]["n_words"] for i in order]), want.stats[0]["n_words"])
    assert sum(tot.per_step_failed.values()) == 900 - tot.kept

    class Boom:
        def run_packed(self, u, o):
            raise RuntimeError("device lost")

    with pytest.raises(RuntimeError, match="device lost"):
        StreamingRunner(lambda: Boom(), pipelines=2).run(RowGroup(i, utf8[:10], off[:2]) for i in range(4))
