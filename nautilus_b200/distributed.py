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
    """nautilus_csm_partition (pure host code of the library, no GPU needed): sort
    pairs by target, cut into world_size contiguous blocks of near-equal total weight.
    -> (order, bounds): order = permutation of pair indices, rank r owns
    order[bounds[r]:bounds[r+1]]."""
    from . import _lib
    b = np.ascontiguousarray(b_ids, dtype=np.int32)
    n = len(b)
    w = np.ascontiguousarray(weights, dtype=np.float64) if n else np.zeros(0)
    order = np.empty(n, np.int32)
    bounds = np.zeros(world_size + 1, np.int64)
    _lib.check(_lib.lib().nautilus_csm_partition(_lib.ptr(b), _lib.ptr(w) if n else None, n,
                                                 world_size, _lib.ptr(order), _lib.ptr(bounds)))
    return order.astype(np.int64), bounds


def pair_weights(n_rot, n_t: int, n_points):
    """Low-res work of a pair: rotations x translations x points."""
    return np.asarray(n_rot, np.float64) * float(n_t * n_t) * np.asarray(n_points, np.float64)


def pair_weights_c(params, a_ids, rot_min, rot_max, offsets):
    """nautilus_csm_pair_weights: the weights the library itself partitions by."""
    import ctypes as C
    from . import _lib
    a = np.ascontiguousarray(a_ids, dtype=np.int32)
    n = len(a)
    rmin = np.ascontiguousarray(np.broadcast_to(np.asarray(rot_min, np.float64), (n,)))
    rmax = np.ascontiguousarray(np.broadcast_to(np.asarray(rot_max, np.float64), (n,)))
    offs = np.ascontiguousarray(offsets, dtype=np.int64)
    w = np.empty(n, np.float64)
    _lib.check(_lib.lib().nautilus_csm_pair_weights(C.byref(params), _lib.ptr(a), _lib.ptr(rmin),
                                                    _lib.ptr(rmax), _lib.ptr(offs), n, _lib.ptr(w)))
    return w


def match_batch_sharded(match_fn, a_ids, b_ids, rot_min, rot_max, weights, *, rank: int,
                        world_size: int, all_gather):
    """Each rank matches its block with match_fn(a, b, rmin, rmax) -> RECORD array
    (device or host), all_gather(local, counts) returns every rank's records in
    rank order; the result is scattered back to caller order.

    The GPU path does all of this inside the library
    (nautilus_csm_match_batch_sharded: same partition code, ncclAllGather); this
    host-side form exists so that the partition / scatter logic also runs without a
    GPU, over gloo, with an injected matcher (tests/test_multi_rank.py)."""
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
