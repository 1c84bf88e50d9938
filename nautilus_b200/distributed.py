"""Multi-GPU plumbing of the batch path: one process per GPU, pairs partitioned
across ranks, only the per-pair result records exchanged.

Pairs are independent (CorrelativeScanMatcher.cc:145-245 keeps no cross-pair
state), so the data path has no collective; the single exchange is an
all_gather of the 48-byte records nautilus_csm_copy_results_dev produces.
The partition keeps every pair that shares a target on one rank (its lookup
table is built once there) unless a target's pairs straddle a block boundary.
"""
from __future__ import annotations

import numpy as np

RECORD = np.dtype([("score", "<f8"), ("x", "<f4"), ("y", "<f4"), ("theta", "<f4"),
                   ("status", "<i4"), ("n_low", "<i8"), ("n_high", "<i8"), ("n_pops", "<i8")])
assert RECORD.itemsize == 48


def partition_pairs(b_ids, weights, world_size: int):
    """Sort pairs by target, cut into world_size contiguous blocks of near-equal
    total weight.  -> (order, bounds): order = permutation of pair indices,
    rank r owns order[bounds[r]:bounds[r+1]]."""
    b_ids = np.asarray(b_ids)
    n = len(b_ids)
    order = np.argsort(b_ids, kind="stable")
    w = np.asarray(weights, dtype=np.float64)[order] if n else np.zeros(0)
    cum = np.cumsum(w)
    total = cum[-1] if n else 0.0
    bounds = [0]
    for r in range(1, world_size):
        bounds.append(int(np.searchsorted(cum, total * r / world_size, side="left")) if n else 0)
    bounds.append(n)
    bounds = np.maximum.accumulate(np.asarray(bounds, dtype=np.int64))
    return order, bounds


def pair_weights(n_rot, n_t: int, n_points):
    """Low-res work of a pair: rotations x translations x points."""
    return np.asarray(n_rot, np.float64) * float(n_t * n_t) * np.asarray(n_points, np.float64)


def match_batch_sharded(match_fn, a_ids, b_ids, rot_min, rot_max, weights, *, rank: int,
                        world_size: int, all_gather):
    """Each rank matches its block with match_fn(a, b, rmin, rmax) -> RECORD array
    (device or host), all_gather(local, counts) returns every rank's records in
    rank order; the result is scattered back to caller order.

    match_fn / all_gather are injected so the same logic runs over NCCL on GPUs
    (bench.py) and over gloo on CPUs (tests/test_multi_rank.py)."""
    a_ids, b_ids = np.asarray(a_ids), np.asarray(b_ids)
    n = len(a_ids)
    rot_min = np.broadcast_to(np.asarray(rot_min, np.float64), (n,))
    rot_max = np.broadcast_to(np.asarray(rot_max, np.float64), (n,))
    order, bounds = partition_pairs(b_ids, weights, world_size)
    mine = order[bounds[rank]:bounds[rank + 1]]
    local = match_fn(a_ids[mine], b_ids[mine], rot_min[mine], rot_max[mine])
    counts = (bounds[1:] - bounds[:-1]).astype(np.int64)
    gathered = all_gather(local, counts)            # RECORD array, rank blocks concatenated
    out = np.empty(n, dtype=RECORD)
    out[order] = gathered
    return out
