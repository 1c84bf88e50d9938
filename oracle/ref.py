"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/_ref/libnautilus_ref.so,
the UNMODIFIED reference CSM compiled by oracle/build.py (see ref_wrap.cc).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this.  The product path
(nautilus_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libnautilus_ref.so")

_lib = None
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(LIB_PATH)
    d, f, vp, i64, u64 = C.c_double, C.c_float, C.c_void_p, C.c_int64, C.c_uint64
    L.ref_table_new.restype = vp
    L.ref_table_new.argtypes = [d, d]
    L.ref_table_free.argtypes = [vp]
    L.ref_table_width.restype = u64
    L.ref_table_width.argtypes = [vp]
    L.ref_table_height.restype = u64
    L.ref_table_height.argtypes = [vp]
    L.ref_table_set.argtypes = [vp, f, f, d]
    L.ref_table_get.restype = d
    L.ref_table_get.argtypes = [vp, f, f]
    L.ref_table_is_inside.restype = C.c_int
    L.ref_table_is_inside.argtypes = [vp, f, f]
    L.ref_table_max_area.restype = d
    L.ref_table_max_area.argtypes = [vp, d, d, d, d]
    L.ref_table_abs_coords.restype = u64
    L.ref_table_abs_coords.argtypes = [vp, d, d]
    L.ref_table_values.argtypes = [vp, _f64p]
    L.ref_table_load.argtypes = [vp, _f64p]
    L.ref_table_blur.argtypes = [vp]
    L.ref_table_normalize.argtypes = [vp]
    L.ref_lowres_from_table.restype = vp
    L.ref_lowres_from_table.argtypes = [d, d, d, d, vp]
    L.ref_highres_table.restype = vp
    L.ref_highres_table.argtypes = [d, d, d, d, _f32p, i64]
    if hasattr(L, "ref_csm_laser_to_points"):  # front end (ref_frontend_wrap.cc)
        L.ref_csm_laser_to_points.restype = i64
        L.ref_csm_laser_to_points.argtypes = [_f32p, i64, f, f, f, f, d, C.c_int, _f32p]
    if hasattr(L, "ref_nautilus_laser_to_points"):  # nautilus front end (ref_nautilus_wrap.cc)
        L.ref_nautilus_laser_to_points.restype = i64
        L.ref_nautilus_laser_to_points.argtypes = [_f32p, i64, f, f, f, f, d, _f32p]
        L.ref_scatter_score.restype = d
        L.ref_scatter_score.argtypes = [_f32p, i64]
        L.ref_lc_candidates.restype = i64
        L.ref_lc_candidates.argtypes = [_f32p, _i64p, C.c_int32, _f64p, _i64p]
    if hasattr(L, "ref_residuals_eval"):  # residual functors (ref_residual_wrap.cc)
        L.ref_residuals_eval.restype = C.c_int
        L.ref_residuals_eval.argtypes = [C.c_int, _f64p, _f64p, _i64p, C.c_int32, _f32p, _f32p,
                                         _f32p, _f32p, _f64p, vp, vp]
    if hasattr(L, "ref_kd_new"):  # KD-tree part (ref_nn_wrap.cc)
        L.ref_kd_new.restype = vp
        L.ref_kd_new.argtypes = [_f32p, i64]
        L.ref_kd_free.argtypes = [vp]
        L.ref_kd_nearest.argtypes = [vp, _f32p, i64, f, _f32p, _i32p]
        L.ref_nn_match_batch.restype = d
        L.ref_nn_match_batch.argtypes = [_f32p, _i64p, C.c_int32, _i32p, _i32p, _f32p, d,
                                         C.c_int32, _i64p, _i32p, vp, C.c_int32]
    pd, pf = C.POINTER(d), C.POINTER(f)
    L.ref_match.argtypes = [d, d, d, d, _f32p, i64, _f32p, i64, d, d, pd, pf, pf, pf]
    L.ref_match_restricted.argtypes = [d, d, d, d, _f32p, i64, _f32p, i64, d, d, d,
                                       pd, pf, pf, pf]
    L.ref_uncertainty2.argtypes = [d, d, d, d, _f32p, i64, _f32p, i64, d, _f32p]
    L.ref_uncertainty3.argtypes = [d, d, d, d, _f32p, i64, _f32p, i64, _f32p]
    L.ref_match_batch.restype = d
    L.ref_match_batch.argtypes = [d, d, d, d, _f32p, _i64p, _i32p, _i32p, _f64p, _f64p,
                                  i64, C.c_int, _f64p, _f32p, _f32p, _f32p]
    L.ref_build_tables_batch.restype = d
    L.ref_build_tables_batch.argtypes = [d, d, d, d, _f32p, _i64p, i64, C.c_int, pd]
    L.ref_max_threads.restype = C.c_int
    _lib = L
    return L


class RefTable:
    """struct LookupTable of the reference (CorrelativeScanMatcher.h:27-110)."""

    def __init__(self, range_=None, resolution=None, handle=None):
        self._L = lib()
        self.h = handle if handle is not None else self._L.ref_table_new(range_, resolution)

    def __del__(self):
        try:
            self._L.ref_table_free(self.h)
        except Exception:
            pass

    @property
    def width(self):
        return int(self._L.ref_table_width(self.h))

    @property
    def height(self):
        return int(self._L.ref_table_height(self.h))

    def set(self, x, y, v):
        self._L.ref_table_set(self.h, x, y, v)

    def get(self, x, y):
        return float(self._L.ref_table_get(self.h, x, y))

    def is_inside(self, x, y):
        return bool(self._L.ref_table_is_inside(self.h, x, y))

    def max_area(self, sx, sy, ex, ey):
        return float(self._L.ref_table_max_area(self.h, sx, sy, ex, ey))

    def abs_coords(self, x, y):
        return int(self._L.ref_table_abs_coords(self.h, x, y))

    def values(self) -> np.ndarray:
        out = np.empty((self.height, self.width), dtype=np.float64)
        self._L.ref_table_values(self.h, out)
        return out

    def load(self, arr):
        arr = np.ascontiguousarray(arr, dtype=np.float64)
        assert arr.shape == (self.height, self.width)
        self._L.ref_table_load(self.h, arr)

    def blur(self):
        self._L.ref_table_blur(self.h)

    def normalize(self):
        self._L.ref_table_normalize(self.h)


class RefMatcher:
    """class CorrelativeScanMatcher of the reference (CorrelativeScanMatcher.h:112-152)."""

    def __init__(self, scanner_range, trans_range, low_res, high_res):
        self._L = lib()
        self.p = (float(scanner_range), float(trans_range), float(low_res), float(high_res))

    def highres(self, cloud) -> RefTable:
        cloud = np.ascontiguousarray(cloud, dtype=np.float32).reshape(-1, 2)
        return RefTable(handle=self._L.ref_highres_table(*self.p, cloud, len(cloud)))

    def lowres(self, hi: RefTable) -> RefTable:
        return RefTable(handle=self._L.ref_lowres_from_table(*self.p, hi.h))

    def match(self, a, b, rot_min=0.0, rot_max=2 * np.pi):
        a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 2)
        b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1, 2)
        s, x, y, t = C.c_double(), C.c_float(), C.c_float(), C.c_float()
        self._L.ref_match(*self.p, a, len(a), b, len(b), rot_min, rot_max,
                          C.byref(s), C.byref(x), C.byref(y), C.byref(t))
        return s.value, np.float32(x.value), np.float32(y.value), np.float32(t.value)

    def match_restricted(self, a, b, rot_a, rot_b, restriction):
        a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 2)
        b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1, 2)
        s, x, y, t = C.c_double(), C.c_float(), C.c_float(), C.c_float()
        self._L.ref_match_restricted(*self.p, a, len(a), b, len(b), rot_a, rot_b,
                                     restriction, C.byref(s), C.byref(x), C.byref(y),
                                     C.byref(t))
        return s.value, np.float32(x.value), np.float32(y.value), np.float32(t.value)

    def uncertainty2(self, a, b, rotation):
        a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 2)
        b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1, 2)
        out = np.zeros(4, dtype=np.float32)
        self._L.ref_uncertainty2(*self.p, a, len(a), b, len(b), rotation, out)
        return out.reshape(2, 2)

    def uncertainty3(self, a, b):
        a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 2)
        b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1, 2)
        out = np.zeros(9, dtype=np.float32)
        self._L.ref_uncertainty3(*self.p, a, len(a), b, len(b), out)
        return out.reshape(3, 3)

    def match_batch(self, xy, offsets, pair_a, pair_b, rot_min, rot_max, threads=0):
        n = len(pair_a)
        score = np.zeros(n, np.float64)
        x = np.zeros(n, np.float32)
        y = np.zeros(n, np.float32)
        th = np.zeros(n, np.float32)
        secs = self._L.ref_match_batch(
            *self.p, np.ascontiguousarray(xy, np.float32),
            np.ascontiguousarray(offsets, np.int64),
            np.ascontiguousarray(pair_a, np.int32), np.ascontiguousarray(pair_b, np.int32),
            np.ascontiguousarray(rot_min, np.float64),
            np.ascontiguousarray(rot_max, np.float64), n, threads, score, x, y, th)
        return secs, score, x, y, th

    def build_tables_batch(self, xy, offsets, threads=0):
        cs = C.c_double()
        secs = self._L.ref_build_tables_batch(
            *self.p, np.ascontiguousarray(xy, np.float32),
            np.ascontiguousarray(offsets, np.int64), len(offsets) - 1, threads,
            C.byref(cs))
        return secs, cs.value


def max_threads() -> int:
    return int(lib().ref_max_threads())


class RefKDTree:
    """KDTree<float, 2> of the reference (src/util/kdtree.{h,cc}), unmodified."""

    def __init__(self, points):
        self._L = lib()
        self.points = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 2)
        self._t = self._L.ref_kd_new(self.points, len(self.points))

    def __del__(self):
        try:
            self._L.ref_kd_free(self._t)
        except Exception:
            pass

    def nearest(self, queries, threshold):
        """FindNearestPoint per query -> (float dist, int index of neighbor_node)."""
        q = np.ascontiguousarray(queries, dtype=np.float32).reshape(-1, 2)
        dist = np.zeros(len(q), np.float32)
        idx = np.zeros(len(q), np.int32)
        self._L.ref_kd_nearest(self._t, q, len(q), np.float32(threshold), dist, idx)
        return dist, idx


def nn_match_batch(xy, offsets, src_ids, tgt_ids, transforms, threshold, threads=0,
                   with_dist=False):
    """GetPointToPointMatching's search (solver.cc:146-170) for a batch of pairs over
    the compiled reference KD-tree.  -> (seconds, match_idx, out_offsets[, dist])."""
    L = lib()
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    offsets = np.ascontiguousarray(offsets, np.int64)
    src_ids = np.ascontiguousarray(src_ids, np.int32)
    tgt_ids = np.ascontiguousarray(tgt_ids, np.int32)
    m = np.ascontiguousarray(transforms, np.float32).reshape(-1, 6)
    sizes = (offsets[1:] - offsets[:-1])[src_ids]
    out_off = np.zeros(len(src_ids) + 1, np.int64)
    np.cumsum(sizes, out=out_off[1:])
    idx = np.full(int(out_off[-1]), -2, np.int32)
    dist = np.zeros(int(out_off[-1]), np.float32) if with_dist else None
    secs = L.ref_nn_match_batch(xy, offsets, len(offsets) - 1, src_ids, tgt_ids, m,
                                float(threshold), len(src_ids), out_off, idx,
                                dist.ctypes.data_as(C.c_void_p) if with_dist else None,
                                threads or max_threads())
    return (secs, idx, out_off, dist) if with_dist else (secs, idx, out_off)


def csm_laser_to_points(ranges, angle_min, angle_max, angle_increment, range_min, max_range,
                        truncate_ends=False):
    """The reference's pointcloud_helpers::LaserScanToPointCloud
    (third_party/csm/src/pointcloud_helpers.cc:44-66), compiled unchanged."""
    r = np.ascontiguousarray(ranges, np.float32)
    out = np.zeros((len(r), 2), np.float32)
    n = lib().ref_csm_laser_to_points(r, len(r), angle_min, angle_max, angle_increment, range_min,
                                      max_range, int(bool(truncate_ends)), out.reshape(-1))
    return out[:n].copy()


def nautilus_laser_to_points(ranges, angle_min, angle_max, angle_increment, range_min, max_range):
    """The reference's nautilus::LaserScanToPointCloud (src/input/pointcloud_helpers.cc:28-48,
    the variant the Nautilus node calls), compiled unchanged."""
    r = np.ascontiguousarray(ranges, np.float32)
    out = np.zeros((len(r), 2), np.float32)
    n = lib().ref_nautilus_laser_to_points(r, len(r), angle_min, angle_max, angle_increment,
                                           range_min, max_range, out.reshape(-1))
    return out[:n].copy()


def scatter_score(cloud) -> float:
    """ComputeScatterMatrixScore (src/loop_closure/lc_candidate_filter.cc:34-52), compiled
    unchanged; the 2x2 eigenvalues come from oracle/shim's closed form (Eigen absent)."""
    c = np.ascontiguousarray(cloud, np.float32).reshape(-1, 2)
    return float(lib().ref_scatter_score(c.reshape(-1), len(c)))


def lc_candidates(xy, offsets, poses):
    """LCCandidateFilter::GetLCCandidates (lc_candidate_filter.cc:66-81), compiled unchanged,
    over a SLAMState2D made of the given clouds and (x, y, theta) poses."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    offsets = np.ascontiguousarray(offsets, np.int64)
    poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 3)
    out = np.zeros(len(offsets) - 1, np.int64)
    n = lib().ref_lc_candidates(xy.reshape(-1), offsets, len(offsets) - 1, poses.reshape(-1), out)
    return out[:n].copy()


def residuals_eval(kind, source_poses, target_poses, offsets, sp, tp, sn=None, tn=None,
                   jacobians=True):
    """The reference's LIDARPointResidual (kind 0) / LIDARNormalResidual (kind 1) functors
    (src/optimization/slam_residuals.h:64-177), compiled unchanged and evaluated through
    AutoDiffCostFunction::Evaluate.  -> (residuals [n, 2], jac_source [n, 2, 3],
    jac_target [n, 2, 3]) like nautilus_residuals_eval."""
    sp = np.ascontiguousarray(sp, np.float32).reshape(-1, 2)
    tp = np.ascontiguousarray(tp, np.float32).reshape(-1, 2)
    n = len(sp)
    sn = np.zeros((n, 2), np.float32) if sn is None else np.ascontiguousarray(sn, np.float32).reshape(-1, 2)
    tn = np.zeros((n, 2), np.float32) if tn is None else np.ascontiguousarray(tn, np.float32).reshape(-1, 2)
    offsets = np.ascontiguousarray(offsets, np.int64)
    a = np.ascontiguousarray(source_poses, np.float64).reshape(-1, 3)
    b = np.ascontiguousarray(target_poses, np.float64).reshape(-1, 3)
    res = np.zeros((n, 2), np.float64)
    js = np.zeros((n, 2, 3), np.float64)
    jt = np.zeros((n, 2, 3), np.float64)
    rc = lib().ref_residuals_eval(kind, a.reshape(-1), b.reshape(-1), offsets, len(offsets) - 1,
                                  sp.reshape(-1), tp.reshape(-1), sn.reshape(-1), tn.reshape(-1),
                                  res.reshape(-1),
                                  js.ctypes.data_as(C.c_void_p) if jacobians else None,
                                  jt.ctypes.data_as(C.c_void_p) if jacobians else None)
    assert rc == 0
    return (res, js, jt) if jacobians else res
