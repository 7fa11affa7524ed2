"""Multi-GPU sharding for the filter chain: documents are independent, so row groups are dealt to
ranks and every rank runs the whole chain on its own GPU.  No data-path collective; the only
communication is a host-side gather of the per-row-group keep masks onto rank 0
(the reference's equivalent is N competing AMQP consumers, src/worker_logic.rs:241-283).
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import numpy as np


def row_groups(n_docs: int, group_docs: int = 65536) -> List[Tuple[int, int]]:
    """[start, end) document ranges of the row groups of a file with n_docs rows."""
    return [(a, min(a + group_docs, n_docs)) for a in range(0, n_docs, group_docs)]


def groups_of_rank(n_groups: int, rank: int, world: int, sizes: Sequence[int] | None = None) -> List[int]:
    """Round-robin by default; with byte sizes, greedy least-loaded (longest first) so that long-document
    row groups balance (deterministic: every rank computes the same assignment)."""
    if sizes is None:
        return list(range(rank, n_groups, world))
    order = sorted(range(n_groups), key=lambda g: (-int(sizes[g]), g))
    load = [0] * world
    mine = []
    for g in order:
        r = min(range(world), key=lambda k: (load[k], k))
        load[r] += int(sizes[g])
        if r == rank:
            mine.append(g)
    return sorted(mine)


def gather_keep_masks(local: Dict[int, np.ndarray], n_groups: int, rank: int, world: int):
    """Gather {row group -> keep mask} from every rank; rank 0 returns the masks ordered by row group
    (others return None).  Uses the default process group (gloo on CPU, nccl-capable object gather)."""
    import torch.distributed as dist

    if world == 1:
        return [local[g] for g in range(n_groups)]
    payload = {int(g): np.packbits(m.astype(np.uint8)) for g, m in local.items()}
    lens = {int(g): int(len(m)) for g, m in local.items()}
    out = [None] * world if rank == 0 else None
    dist.gather_object((payload, lens), out, dst=0)
    if rank != 0:
        return None
    masks: List[np.ndarray] = [None] * n_groups  # type: ignore
    for payload_r, lens_r in out:
        for g, bits in payload_r.items():
            masks[g] = np.unpackbits(bits)[: lens_r[g]].astype(np.uint8)
    assert all(m is not None for m in masks), "a row group was not processed by any rank"
    return masks
