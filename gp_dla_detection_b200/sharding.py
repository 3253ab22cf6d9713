"""Quasar sharding across the GPUs of one box.

Every quasar is an independent unit, so the path shards by quasar with NO data-path collective:
a static longest-processing-time partition with cost n_q * S, each rank evaluates its shard on its
own GPU through its own dla_ctx, and the results are concatenated on the host ("final host gather").
torch.distributed is plumbing only (barrier, timing reduction, the host gather).
"""
from __future__ import annotations

import numpy as np


def partition_quasars(costs, world_size: int):
    """Static LPT partition. Returns a list (one entry per rank) of index arrays into the batch,
    each sorted ascending (stable output order inside a shard)."""
    costs = np.asarray(costs, dtype=np.float64)
    order = np.argsort(-costs, kind="stable")
    loads = np.zeros(world_size)
    shards = [[] for _ in range(world_size)]
    for q in order:
        r = int(np.argmin(loads))
        shards[r].append(int(q))
        loads[r] += costs[q]
    return [np.array(sorted(s), dtype=np.int64) for s in shards]


def gather_outputs(local_out: dict, local_index: np.ndarray, num_quasars: int, dist=None, dst: int = 0):
    """Concatenate per-rank output dicts (arrays with leading dimension = shard size) back into batch
    order on rank `dst`.  dist=None: single process (returns a reordered copy)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        pieces = [(local_index, local_out)]
    else:
        gathered = [None] * dist.get_world_size() if dist.get_rank() == dst else None
        dist.gather_object((local_index, local_out), gathered, dst=dst)
        if dist.get_rank() != dst:
            return None
        pieces = gathered
    merged = {}
    for key, val in pieces[0][1].items():
        if val is None:
            merged[key] = None
            continue
        proto = np.asarray(val)
        fill = np.nan if proto.dtype.kind == "f" else -1
        merged[key] = np.full((num_quasars,) + proto.shape[1:], fill, dtype=proto.dtype)
    for idx, out in pieces:
        for key, val in out.items():
            if val is not None and len(idx):
                merged[key][idx] = val
    return merged
