"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy float32, libm sinf/cosf) of
the front end that feeds the scan-matching path:

  nautilus::LaserScanToPointCloud            src/input/pointcloud_helpers.cc:28-48
  pointcloud_helpers::LaserScanToPointCloud  third_party/csm/src/pointcloud_helpers.cc:44-66
  FilterGlancing                             third_party/csm/src/pointcloud_helpers.cc:21-42
  ComputeMean / ComputeScatterMatrixScore    src/loop_closure/lc_candidate_filter.cc:23-52
  DistantFromLastScan / GetLCCandidates      src/loop_closure/lc_candidate_filter.cc:54-81

The csm variant is pinned against the reference's own translation unit compiled
unchanged (oracle/ref.py: csm_laser_to_points).  The nautilus variant shares its
arithmetic (Rotation2Df(angle) * (range, 0)) and differs only in how the angle
is formed; lc_candidate_filter.cc needs ROS and Eigen::EigenSolver (Eigen 3.3.7,
un-vendored): PARITY UNPINNED for the scatter score beyond the closed-form
eigenvalues of the symmetric 2x2 matrix (same closed form as oracle/shim).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.
"""
from __future__ import annotations

import ctypes as C
import ctypes.util

import numpy as np

_libm = C.CDLL(ctypes.util.find_library("m") or "libm.so.6")
_libm.sinf.restype = C.c_float
_libm.sinf.argtypes = [C.c_float]
_libm.cosf.restype = C.c_float
_libm.cosf.argtypes = [C.c_float]
F = np.float32


def _rotate_range(angles: np.ndarray, ranges: np.ndarray) -> np.ndarray:
    """Rotation2Df(angle).toRotationMatrix() * Vector2f(range, 0): each component is
    m(i,0)*range + m(i,1)*0 with separately rounded products (no FMA)."""
    c = np.array([_libm.cosf(float(a)) for a in angles], F)
    s = np.array([_libm.sinf(float(a)) for a in angles], F)
    zero = F(0.0)
    x = c * ranges + (-s) * zero
    y = s * ranges + c * zero
    return np.stack([x, y], axis=1).astype(F)


def laser_to_points_nautilus(ranges, angle_min, angle_increment, range_min, max_range):
    """src/input/pointcloud_helpers.cc:28-48: angle = angle_min + increment * index (float)."""
    ranges = np.asarray(ranges, F)
    idx = np.arange(len(ranges)).astype(F)               # size_t -> float
    angles = (F(angle_min) + F(angle_increment) * idx).astype(F)
    with np.errstate(invalid="ignore"):
        ok = (ranges >= F(range_min)) & (ranges.astype(np.float64) <= float(max_range))
    return _rotate_range(angles[ok], ranges[ok])


def csm_beam_angles(n, angle_min, angle_increment):
    """angle_offset accumulated in float (third_party/csm/src/pointcloud_helpers.cc:48,63)."""
    out = np.empty(n, F)
    a = F(angle_min)
    inc = F(angle_increment)
    for i in range(n):
        out[i] = a
        a = F(a + inc)
    return out


def laser_to_points_csm(ranges, angle_min, angle_max, angle_increment, range_min, max_range,
                        truncate_ends=False):
    """third_party/csm/src/pointcloud_helpers.cc:44-66 followed by FilterGlancing."""
    ranges = np.asarray(ranges, F)
    angles = csm_beam_angles(len(ranges), angle_min, angle_increment)
    trunc = F(np.pi / 12.0) if truncate_ends else F(0)
    lo = F(F(angle_min) + trunc)
    hi = F(F(angle_max) - trunc)
    with np.errstate(invalid="ignore"):
        ok = (ranges >= F(range_min)) & (ranges.astype(np.float64) <= float(max_range))
    ok &= (angles > lo) & (angles < hi)
    pts = _rotate_range(angles[ok], ranges[ok])
    if len(pts) == 0:
        return pts
    prev = np.vstack([pts[:1], pts[:-1]])
    d = prev - pts
    norm = np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(F)
    keep = norm.astype(np.float64) <= 0.10
    return pts[keep]


def scatter_score(cloud) -> float:
    """lc_candidate_filter.cc:23-52.  Float accumulation in point order; the double
    factor 1.0 / size is narrowed to float before it scales the sum (Eigen's
    promote_scalar_arg); eigenvalues of the 2x2 scatter matrix in closed form."""
    p = np.asarray(cloud, F).reshape(-1, 2)
    n = len(p)
    if n == 0:
        return float("nan")
    s = np.cumsum(p, axis=0, dtype=F)[-1]                # sequential float sums
    mean = (F(1.0 / n) * s).astype(F)
    d = (p - mean).astype(F)
    m00 = np.cumsum(d[:, 0] * d[:, 0], dtype=F)[-1]
    m01 = np.cumsum(d[:, 0] * d[:, 1], dtype=F)[-1]
    m10 = np.cumsum(d[:, 1] * d[:, 0], dtype=F)[-1]
    m11 = np.cumsum(d[:, 1] * d[:, 1], dtype=F)[-1]
    a, b, c, dd = float(m00), float(m01), float(m10), float(m11)
    tr, det = a + dd, a * dd - b * c
    disc = tr * tr / 4.0 - det
    root = np.sqrt(disc) if disc > 0 else 0.0
    e1 = float(F(tr / 2.0 + root))
    e2 = float(F(tr / 2.0 - root))
    lo, hi = min(e1, e2), max(e1, e2)
    with np.errstate(invalid="ignore", divide="ignore"):
        return float(np.float64(lo) / np.float64(hi))


def lc_candidates(pose_xy, scores, min_dist=5.0, min_score=0.70):
    """GetLCCandidates (lc_candidate_filter.cc:66-81): a node is a keyframe candidate
    when it is at least min_dist from the last accepted one (float norm, compared in
    double) and its scatter score is >= min_score."""
    pose_xy = np.asarray(pose_xy, F).reshape(-1, 2)
    out = []
    for i in range(len(pose_xy)):
        if out:
            d = pose_xy[i] - pose_xy[out[-1]]
            if float(np.sqrt(F(d[0] * d[0] + d[1] * d[1]))) < min_dist:
                continue
        if scores[i] >= min_score:
            out.append(i)
    return np.asarray(out, np.int64)
