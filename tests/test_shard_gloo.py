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
    pl = oracle.OraclePipeline(CHAINS["baseline_order"])
    local = {}
    for g in mine:
        a, b = groups[g]
        local[g] = pl.run_packed(utf8[off[a]:off[b]], off[a:b + 1] - off[a]).keep
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
