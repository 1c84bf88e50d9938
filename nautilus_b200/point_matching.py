"""Host-side mirror of the reference's point-to-point matching interface
(src/optimization/solver.cc:67-172, src/optimization/data_structures.h:62-100)
over the C ABI.  The search runs on the B200 (nautilus_nn_match_batch); this
module only assembles PointCorrespondences from the returned indices, the way
Solver::GetPointToPointMatching fills its result."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import ptr


def pose_transform(source_pose, target_pose) -> np.ndarray:
    """target_to_world.inverse() * source_to_world as a row-major 2x3 float matrix
    (slam_util.h:20-40, solver.cc:148-159)."""
    s = np.ascontiguousarray(source_pose, dtype=np.float64)
    t = np.ascontiguousarray(target_pose, dtype=np.float64)
    m = np.empty(6, np.float32)
    _lib.lib().nautilus_nn_pose_transform(ptr(s), ptr(t), ptr(m))
    return m


@dataclass
class PointCorrespondences:
    """struct PointCorrespondences (data_structures.h:62-100)."""
    source_index: int
    target_index: int
    source_pose: np.ndarray | None = None
    target_pose: np.ndarray | None = None
    source_points: np.ndarray = field(default_factory=lambda: np.zeros((0, 2), np.float32))
    target_points: np.ndarray = field(default_factory=lambda: np.zeros((0, 2), np.float32))
    source_normals: np.ndarray = field(default_factory=lambda: np.zeros((0, 2), np.float32))
    target_normals: np.ndarray = field(default_factory=lambda: np.zeros((0, 2), np.float32))
    difference: float = 0.0


def correspondences_from_matches(matcher, xy, offsets, normals, poses, src_ids, tgt_ids,
                                 match_idx, out_offsets, self_index=None):
    """AddCorrespondence (data_structures.h:89-99) for every matched source point,
    in source order.  `normals` has one row per uploaded point; the normal of a
    point is read through self_index, as GetPointNormal's tree lookup does."""
    xy = np.asarray(xy, np.float32).reshape(-1, 2)
    normals = np.asarray(normals, np.float32).reshape(-1, 2)
    if self_index is None:
        self_index = matcher.nn_self_index()
    out = []
    for p, (s, t) in enumerate(zip(src_ids, tgt_ids)):
        m = match_idx[out_offsets[p]:out_offsets[p + 1]]
        hit = np.nonzero(m >= 0)[0]
        so, to = int(offsets[s]), int(offsets[t])
        sp = xy[so + hit]
        tp = xy[to + m[hit]]
        c = PointCorrespondences(int(s), int(t), None if poses is None else poses[s],
                                 None if poses is None else poses[t])
        c.source_points, c.target_points = sp, tp
        c.source_normals = normals[so + self_index[so + hit]]
        c.target_normals = normals[to + self_index[to + m[hit]]]
        # difference += (source_point - target_point).norm(): float norm, double sum, in order
        d = sp - tp
        nrm = np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(np.float32)).astype(np.float32)
        acc = 0.0
        for v in nrm:
            acc += float(v)
        c.difference = acc
        out.append(c)
    return out


def GetPointToPointMatching(matcher, xy, offsets, normals, poses, source_node_idx,
                            target_node_idx, outlier_threshold):
    """Solver::GetPointToPointMatching (solver.cc:132-172) for one pair, type ALL."""
    m = pose_transform(poses[source_node_idx], poses[target_node_idx])
    idx, offs = matcher.nn_match_batch([source_node_idx], [target_node_idx], m, outlier_threshold)
    return correspondences_from_matches(matcher, xy, offsets, normals, poses,
                                        [source_node_idx], [target_node_idx], idx, offs)[0]
