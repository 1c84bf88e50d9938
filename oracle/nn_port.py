"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/_build/libnn_oracle.so, the
CPU restatement of the reference's nearest-neighbour path (nn_oracle.cc).  Only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import build

_lib = None
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def lib():
    global _lib
    if _lib is None:
        build.build_port()
        L = C.CDLL(build.NN_PORT_LIB)
        L.nn_port_tree_new.restype = C.c_void_p
        L.nn_port_tree_new.argtypes = [_f32p, C.c_int64]
        L.nn_port_tree_free.argtypes = [C.c_void_p]
        L.nn_port_match.argtypes = [C.c_void_p, C.c_int64, _f32p, C.c_int64, _f32p, C.c_double,
                                    _i32p, _f32p]
        _lib = L
    return _lib


def match_batch(xy, offsets, src_ids, tgt_ids, transforms, threshold):
    """GetPointToPointMatching's search for a batch of pairs -> (idx, out_offsets, dist)."""
    L = lib()
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    m = np.ascontiguousarray(transforms, np.float32).reshape(-1, 6)
    sizes = (np.asarray(offsets[1:]) - np.asarray(offsets[:-1]))[np.asarray(src_ids)]
    out_off = np.zeros(len(src_ids) + 1, np.int64)
    np.cumsum(sizes, out=out_off[1:])
    idx = np.full(int(out_off[-1]), -2, np.int32)
    dist = np.zeros(int(out_off[-1]), np.float32)
    trees = {}
    for p, (s, t) in enumerate(zip(src_ids, tgt_ids)):
        s, t = int(s), int(t)
        if t not in trees:
            c = np.ascontiguousarray(xy[offsets[t]:offsets[t + 1]])
            trees[t] = (L.nn_port_tree_new(c, len(c)), len(c))
        src = np.ascontiguousarray(xy[offsets[s]:offsets[s + 1]])
        n = len(src)
        i_out = np.zeros(max(n, 1), np.int32)
        d_out = np.zeros(max(n, 1), np.float32)
        L.nn_port_match(trees[t][0], trees[t][1], src, n, np.ascontiguousarray(m[p]),
                        float(threshold), i_out, d_out)
        idx[out_off[p]:out_off[p + 1]] = i_out[:n]
        dist[out_off[p]:out_off[p + 1]] = d_out[:n]
    for t, _ in trees.values():
        L.nn_port_tree_free(t)
    return idx, out_off, dist


def brute_force(xy, offsets, src_ids, tgt_ids, transforms, threshold):
    """Exact nearest neighbour in float (independent check; differs from the tree
    where the reference's walk skips a single-child leaf, see nn_kernels.cu)."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    m = np.ascontiguousarray(transforms, np.float32).reshape(-1, 6)
    out = []
    for p, (s, t) in enumerate(zip(src_ids, tgt_ids)):
        src = xy[offsets[s]:offsets[s + 1]]
        tgt = xy[offsets[t]:offsets[t + 1]]
        q = np.stack([(m[p, 0] * src[:, 0] + m[p, 1] * src[:, 1]) + m[p, 2],
                      (m[p, 3] * src[:, 0] + m[p, 4] * src[:, 1]) + m[p, 5]], 1).astype(np.float32)
        if len(tgt) == 0:
            out.append(np.full(len(src), -1, np.int32))
            continue
        d = tgt[None, :, :] - q[:, None, :]
        dist = np.sqrt((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]).astype(np.float32))
        i = dist.argmin(1).astype(np.int32)
        best = dist[np.arange(len(src)), i]
        out.append(np.where(best.astype(np.float64) < threshold, i, -1).astype(np.int32))
    return np.concatenate(out) if out else np.zeros(0, np.int32)
