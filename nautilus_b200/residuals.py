"""Host-side mirror of the lidar residual functors over the C ABI
(src/optimization/slam_residuals.h:64-177): batched evaluation of
LIDARPointResidual / LIDARNormalResidual blocks with analytic Jacobians on the
device.  No CPU fallback."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import check, pinned_empty, ptr

POINT, NORMAL = 0, 1


def evaluate_residual_blocks(matcher, kind, source_poses, target_poses, offsets, source_points,
                             target_points, source_normals=None, target_normals=None,
                             jacobians: bool = True):
    """-> residuals [n, 2] and, with jacobians, (jac_source [n, 2, 3], jac_target [n, 2, 3]):
    a block's rows [offsets[b], offsets[b+1]) flattened are Ceres' residuals /
    jacobians[0] / jacobians[1] of AutoDiffCostFunction<..., DYNAMIC, 3, 3>."""
    sp = np.ascontiguousarray(source_poses, np.float64).reshape(-1, 3)
    tp = np.ascontiguousarray(target_poses, np.float64).reshape(-1, 3)
    offsets = np.ascontiguousarray(offsets, np.int64)
    nb = len(offsets) - 1
    if len(sp) != nb or len(tp) != nb:
        raise ValueError("one source and one target pose per block")
    n = int(offsets[-1]) if nb >= 0 and len(offsets) else 0
    f = lambda a: None if a is None else np.ascontiguousarray(a, np.float32).reshape(-1, 2)  # noqa: E731
    p, q, ns, nt = f(source_points), f(target_points), f(source_normals), f(target_normals)
    res = np.zeros((n, 2), np.float64)
    js = np.zeros((n, 2, 3), np.float64) if jacobians else None
    jt = np.zeros((n, 2, 3), np.float64) if jacobians else None
    o = lambda a: None if a is None else ptr(a)  # noqa: E731
    check(matcher._L.nautilus_residuals_eval(matcher._h, int(kind), ptr(sp), ptr(tp), ptr(offsets), nb,
                                             o(p), o(q), o(ns), o(nt), ptr(res), o(js), o(jt)))
    return (res, js, jt) if jacobians else res


def evaluate_resident_blocks(matcher, kind, source_poses, target_poses, corr_offsets,
                             jacobians: bool = True, pinned: bool = False, out=None):
    """evaluate_residual_blocks over the correspondence set nn_match_resident left on
    the device (one block per pair of that call): only poses go in and residuals /
    Jacobians come out.  Repeatable with new poses, as Ceres re-evaluates a cost function."""
    sp = np.ascontiguousarray(source_poses, np.float64).reshape(-1, 3)
    tp = np.ascontiguousarray(target_poses, np.float64).reshape(-1, 3)
    nb = len(corr_offsets) - 1
    if len(sp) != nb or len(tp) != nb:
        raise ValueError("one source and one target pose per block")
    n = int(corr_offsets[-1])
    if out is not None:          # buffers of an earlier call (a solver iterates on the same set)
        res, js, jt = out
    else:
        new = pinned_empty if pinned else np.zeros   # pinned: D2H at PCIe speed (nautilus_host_alloc)
        res = new((n, 2), np.float64)
        js = new((n, 2, 3), np.float64) if jacobians else None
        jt = new((n, 2, 3), np.float64) if jacobians else None
    o = lambda a: None if a is None else ptr(a)  # noqa: E731
    check(matcher._L.nautilus_residuals_eval_resident(matcher._h, int(kind), ptr(sp), ptr(tp),
                                                      ptr(res), o(js), o(jt)))
    return (res, js, jt) if jacobians else res


def last_kernel_ms(matcher) -> float:
    ms = C.c_float()
    check(matcher._L.nautilus_residuals_last_ms(matcher._h, C.byref(ms)))
    return float(ms.value)
