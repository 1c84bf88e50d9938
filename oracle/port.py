"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/_build/libcsm_oracle.so, the
plain-C restatement of the reference CSM path (csm_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this; nautilus_b200/ never does.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import build as _build

_lib = None
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(_build.build_port())
    d, f, i64, u64 = C.c_double, C.c_float, C.c_int64, C.c_uint64
    pd, pf = C.POINTER(d), C.POINTER(f)
    L.oracle_table_dim.restype = u64
    L.oracle_table_dim.argtypes = [d, d]
    L.oracle_cell.restype = i64
    L.oracle_cell.argtypes = [f, d, u64]
    L.oracle_get_point_value.restype = d
    L.oracle_get_point_value.argtypes = [_f64p, u64, u64, d, f, f]
    L.oracle_set_point_value.argtypes = [_f64p, u64, u64, d, f, f, d]
    L.oracle_max_area.restype = d
    L.oracle_max_area.argtypes = [_f64p, u64, u64, d, d, d, d, d]
    L.oracle_abs_coords.restype = u64
    L.oracle_abs_coords.argtypes = [u64, u64, d, d, d]
    L.oracle_blur.argtypes = [_f64p, u64, u64, d]
    L.oracle_normalize.argtypes = [_f64p, u64]
    L.oracle_highres_table.argtypes = [d, d, _f32p, i64, _f64p]
    L.oracle_lowres_table.argtypes = [d, d, _f64p, u64, u64, d, _f64p]
    L.oracle_rotate.argtypes = [_f32p, i64, d, _f32p]
    L.oracle_cost.restype = d
    L.oracle_cost.argtypes = [_f32p, i64, d, d, _f64p, u64, d]
    L.oracle_search.restype = C.c_int
    L.oracle_search.argtypes = [d, d, d, _f32p, i64, _f64p, u64, _f64p, u64, d, d, d,
                                pd, pf, pf, pf, _i64p]
    L.oracle_match.restype = C.c_int
    L.oracle_match.argtypes = [d, d, d, d, _f32p, i64, _f32p, i64, d, d, d,
                               pd, pf, pf, pf, _i64p]
    L.oracle_rotation_window.argtypes = [d, d, d, pd, pd]
    _lib = L
    return L


ROT_STEP = np.pi / 180  # CorrelativeScanMatcher.cc:92


def _cloud(a):
    return np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 2)


class PortTable:
    """struct LookupTable (CorrelativeScanMatcher.h:27-110) over a numpy array."""

    def __init__(self, range_, resolution, values=None):
        self._L = lib()
        self.resolution = float(resolution)
        self.width = self.height = int(self._L.oracle_table_dim(range_, resolution))
        self.values = (np.zeros((self.height, self.width), np.float64) if values is None
                       else np.ascontiguousarray(values, np.float64))

    def set(self, x, y, v):
        self._L.oracle_set_point_value(self.values, self.width, self.height,
                                       self.resolution, x, y, v)

    def get(self, x, y):
        return float(self._L.oracle_get_point_value(self.values, self.width, self.height,
                                                    self.resolution, x, y))

    def max_area(self, sx, sy, ex, ey):
        return float(self._L.oracle_max_area(self.values, self.width, self.height,
                                             self.resolution, sx, sy, ex, ey))

    def abs_coords(self, x, y):
        return int(self._L.oracle_abs_coords(self.width, self.height, self.resolution, x, y))

    def blur(self, sigma=4.0):
        self._L.oracle_blur(self.values, self.width, self.height, sigma)

    def normalize(self):
        self._L.oracle_normalize(self.values.reshape(-1), self.values.size)


class PortMatcher:
    """class CorrelativeScanMatcher (CorrelativeScanMatcher.h:112-152), restated."""

    def __init__(self, scanner_range, trans_range, low_res, high_res, rot_step=ROT_STEP):
        self._L = lib()
        self.range = float(scanner_range)
        self.trans_range = float(trans_range)
        self.low_res = float(low_res)
        self.high_res = float(high_res)
        self.rot_step = float(rot_step)

    def highres(self, cloud) -> PortTable:
        cloud = _cloud(cloud)
        t = PortTable(self.range, self.high_res)
        self._L.oracle_highres_table(self.range, self.high_res, cloud, len(cloud),
                                     t.values.reshape(-1))
        return t

    def lowres(self, hi: PortTable) -> PortTable:
        t = PortTable(self.range, self.low_res)
        self._L.oracle_lowres_table(self.range, self.low_res, hi.values.reshape(-1),
                                    hi.width, hi.height, hi.resolution,
                                    t.values.reshape(-1))
        return t

    def search(self, a, hi: PortTable, low: PortTable, rot_min=0.0, rot_max=2 * np.pi):
        a = _cloud(a)
        s, x, y, t = C.c_double(), C.c_float(), C.c_float(), C.c_float()
        stats = np.zeros(3, np.int64)
        rc = self._L.oracle_search(self.trans_range, self.low_res, self.high_res, a, len(a),
                                   hi.values.reshape(-1), hi.width, low.values.reshape(-1),
                                   low.width, rot_min, rot_max, self.rot_step,
                                   C.byref(s), C.byref(x), C.byref(y), C.byref(t), stats)
        if rc != 0:
            raise ValueError(f"oracle_search rc={rc}")
        return (s.value, np.float32(x.value), np.float32(y.value), np.float32(t.value)), stats

    def match(self, a, b, rot_min=0.0, rot_max=2 * np.pi, with_stats=False):
        a, b = _cloud(a), _cloud(b)
        s, x, y, t = C.c_double(), C.c_float(), C.c_float(), C.c_float()
        stats = np.zeros(3, np.int64)
        rc = self._L.oracle_match(self.range, self.trans_range, self.low_res, self.high_res,
                                  a, len(a), b, len(b), rot_min, rot_max, self.rot_step,
                                  C.byref(s), C.byref(x), C.byref(y), C.byref(t), stats)
        if rc != 0:
            raise ValueError(f"oracle_match rc={rc}")
        r = (s.value, np.float32(x.value), np.float32(y.value), np.float32(t.value))
        return (r, stats) if with_stats else r

    def rotation_window(self, rot_a, rot_b, restriction):
        lo, hi = C.c_double(), C.c_double()
        self._L.oracle_rotation_window(rot_a, rot_b, restriction, C.byref(lo), C.byref(hi))
        return lo.value, hi.value

    def rotate(self, a, rotation):
        a = _cloud(a)
        out = np.empty_like(a)
        self._L.oracle_rotate(a, len(a), rotation, out)
        return out
