"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the lidar residual functors and
of what ceres::AutoDiffCostFunction makes of them: the functor body is evaluated
literally on forward-mode dual numbers ("jets" with 6 derivative slots: source
pose x, y, theta then target pose x, y, theta), as Ceres 1.14 does.

  LIDARNormalResidual::operator()  src/optimization/slam_residuals.h:64-90
  LIDARPointResidual::operator()   src/optimization/slam_residuals.h:123-145
  PoseArrayToAffine                src/util/slam_util.h:20-28
  Eigen::Transform<T,2,Affine>::inverse(): linear part inverted through the 2x2
  cofactor formula, translation = -(linear^-1 * t)   (Eigen 3.3.7, un-vendored)

Ceres and Eigen are absent from /root/reference and the reference has no test for
the residuals: PARITY UNPINNED (restatement only).  Floating point: tolerance
1e-12 (relative to max(1, |value|)) in the tests.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this."""
from __future__ import annotations

import numpy as np


class Jet:
    """value [n] + derivatives [n, 6]; just enough arithmetic for the functor."""
    __array_ufunc__ = None  # ndarray (op) Jet defers to the reflected method

    def __init__(self, v, d=None):
        self.v = np.asarray(v, np.float64)
        self.d = np.zeros(self.v.shape + (6,)) if d is None else d

    @staticmethod
    def lift(x):
        return x if isinstance(x, Jet) else Jet(np.asarray(x, np.float64))

    def __add__(self, o):
        o = Jet.lift(o)
        return Jet(self.v + o.v, self.d + o.d)

    __radd__ = __add__

    def __sub__(self, o):
        o = Jet.lift(o)
        return Jet(self.v - o.v, self.d - o.d)

    def __rsub__(self, o):
        return Jet.lift(o) - self

    def __neg__(self):
        return Jet(-self.v, -self.d)

    def __mul__(self, o):
        o = Jet.lift(o)
        return Jet(self.v * o.v, self.v[..., None] * o.d + o.v[..., None] * self.d)

    __rmul__ = __mul__

    def __truediv__(self, o):
        o = Jet.lift(o)
        inv = 1.0 / o.v
        q = self.v * inv
        return Jet(q, (self.d - q[..., None] * o.d) * inv[..., None])


def jsin(a):
    return Jet(np.sin(a.v), np.cos(a.v)[..., None] * a.d)


def jcos(a):
    return Jet(np.cos(a.v), -np.sin(a.v)[..., None] * a.d)


def _affine(theta, tx, ty):
    """PoseArrayToAffine: Translation2T(t) * Rotation2DT(theta).toRotationMatrix()."""
    c, s = jcos(theta), jsin(theta)
    return ((c, -s), (s, c)), (tx, ty)


def _inverse(lin, t):
    (a, b), (c, d) = lin
    det = a * d - b * c
    inv = ((d / det, -b / det), (-c / det, a / det))
    ti = (-(inv[0][0] * t[0] + inv[0][1] * t[1]), -(inv[1][0] * t[0] + inv[1][1] * t[1]))
    return inv, ti


def _compose(l1, t1, l2, t2):
    lin = tuple(tuple(l1[i][0] * l2[0][j] + l1[i][1] * l2[1][j] for j in range(2)) for i in range(2))
    t = tuple(l1[i][0] * t2[0] + l1[i][1] * t2[1] + t1[i] for i in range(2))
    return lin, t


def evaluate(kind, source_poses, target_poses, offsets, sp, tp, sn=None, tn=None):
    """-> residuals [n, 2], jac_source [n, 2, 3], jac_target [n, 2, 3]."""
    offsets = np.asarray(offsets, np.int64)
    n = int(offsets[-1])
    blk = np.searchsorted(offsets, np.arange(n), side="right") - 1
    ps = np.asarray(source_poses, np.float64).reshape(-1, 3)[blk]
    pt = np.asarray(target_poses, np.float64).reshape(-1, 3)[blk]

    def var(col, slot):
        d = np.zeros((n, 6))
        d[:, slot] = 1.0
        return Jet(col.copy(), d)

    xs, ys, ths = var(ps[:, 0], 0), var(ps[:, 1], 1), var(ps[:, 2], 2)
    xt, yt, tht = var(pt[:, 0], 3), var(pt[:, 1], 4), var(pt[:, 2], 5)
    s2w = _affine(ths, xs, ys)
    w2t = _inverse(*_affine(tht, xt, yt))
    lin, t = _compose(w2t[0], w2t[1], s2w[0], s2w[1])
    sp = np.asarray(sp, np.float32).reshape(-1, 2).astype(np.float64)
    tp = np.asarray(tp, np.float32).reshape(-1, 2).astype(np.float64)
    px = lin[0][0] * sp[:, 0] + lin[0][1] * sp[:, 1] + t[0]
    py = lin[1][0] * sp[:, 0] + lin[1][1] * sp[:, 1] + t[1]
    if kind == 0:
        r0, r1 = tp[:, 0] - px, tp[:, 1] - py
    else:
        sn = np.asarray(sn, np.float32).reshape(-1, 2).astype(np.float64)
        tn = np.asarray(tn, np.float32).reshape(-1, 2).astype(np.float64)
        r0 = tn[:, 0] * (px - tp[:, 0]) + tn[:, 1] * (py - tp[:, 1])
        r1 = sn[:, 0] * (tp[:, 0] - px) + sn[:, 1] * (tp[:, 1] - py)
    res = np.stack([r0.v, r1.v], axis=1)
    jac = np.stack([r0.d, r1.d], axis=1)            # [n, 2, 6]
    return res, jac[:, :, :3].copy(), jac[:, :, 3:].copy()
