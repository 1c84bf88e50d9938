"""Residual blocks (residuals.cu).  The checker is pinned to the reference's own functor
source: slam_residuals.h:64-177 compiled UNCHANGED into oracle/_ref and evaluated through
AutoDiffCostFunction::Evaluate (fixtures: tests/golden/residual_golden.npz); the numpy
dual-number port (oracle/residual_port.py) must reproduce those fixtures bit for bit and is
itself checked against finite differences.  GPU parity of the analytic-Jacobian kernel
against both.  Tolerance (floating point): 1e-12 relative to max(1, |value|)."""
import os
import numpy as np
import pytest

from conftest import has_gpu
from oracle import residual_port as rp

TOL = 1e-12


def make_blocks(n_blocks, seed=0, max_corr=40):
    rng = np.random.default_rng(seed)
    sizes = rng.integers(0, max_corr, size=n_blocks)
    sizes[0] = max_corr
    offs = np.zeros(n_blocks + 1, np.int64)
    offs[1:] = np.cumsum(sizes)
    n = int(offs[-1])
    sp = rng.uniform(-20, 20, size=(n_blocks, 3)); sp[:, 2] = rng.uniform(-np.pi, np.pi, n_blocks)
    tp = sp + rng.normal(0, 0.3, size=(n_blocks, 3))
    p = rng.uniform(-25, 25, size=(n, 2)).astype(np.float32)
    q = (p + rng.normal(0, 0.1, size=(n, 2))).astype(np.float32)
    ang = rng.uniform(-np.pi, np.pi, size=(2, n))
    ns = np.stack([np.cos(ang[0]), np.sin(ang[0])], 1).astype(np.float32)
    nt = np.stack([np.cos(ang[1]), np.sin(ang[1])], 1).astype(np.float32)
    return sp, tp, offs, p, q, ns, nt


def residual_golden():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "residual_golden.npz"))


@pytest.mark.parametrize("kind", [0, 1])
def test_port_vs_reference_functor_golden(kind):
    g = residual_golden()
    k = f"k{kind}_"
    got = rp.evaluate(kind, g[k + "sp"], g[k + "tp"], g[k + "offs"], g[k + "p"], g[k + "q"],
                      g[k + "ns"], g[k + "nt"])
    for a, name in zip(got, ("res", "js", "jt")):
        assert np.array_equal(a, g[k + name]), name


@pytest.mark.parametrize("kind", [0, 1])
def test_port_vs_compiled_reference_functors(ref_lib, kind):
    sp, tp, offs, p, q, ns, nt = make_blocks(30, seed=21 + kind)
    want = ref_lib.residuals_eval(kind, sp, tp, offs, p, q, ns, nt)
    got = rp.evaluate(kind, sp, tp, offs, p, q, ns, nt)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)
    only = ref_lib.residuals_eval(kind, sp, tp, offs, p, q, ns, nt, jacobians=False)
    assert np.array_equal(only, want[0])


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [0, 1])
def test_gpu_vs_reference_functor_golden(kind):
    from nautilus_b200 import CorrelativeScanMatcher
    from nautilus_b200.residuals import evaluate_residual_blocks
    g = residual_golden()
    k = f"k{kind}_"
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    got = evaluate_residual_blocks(m, kind, g[k + "sp"], g[k + "tp"], g[k + "offs"], g[k + "p"],
                                   g[k + "q"], g[k + "ns"], g[k + "nt"])
    for a, name in zip(got, ("res", "js", "jt")):
        want = g[k + name]
        assert np.all(np.abs(a - want) <= TOL * np.maximum(1.0, np.abs(want))), name
    m.close()


@pytest.mark.parametrize("kind", [0, 1])
def test_oracle_jets_vs_finite_differences(kind):
    sp, tp, offs, p, q, ns, nt = make_blocks(6, seed=1)
    res, js, jt = rp.evaluate(kind, sp, tp, offs, p, q, ns, nt)
    eps = 1e-6
    for which, jac in ((0, js), (1, jt)):
        for c in range(3):
            a, b = [sp.copy(), tp.copy()], [sp.copy(), tp.copy()]
            a[which][:, c] += eps
            b[which][:, c] -= eps
            fd = (rp.evaluate(kind, a[0], a[1], offs, p, q, ns, nt)[0] -
                  rp.evaluate(kind, b[0], b[1], offs, p, q, ns, nt)[0]) / (2 * eps)
            assert np.allclose(jac[:, :, c], fd, rtol=0, atol=2e-7)


def test_oracle_identity_pose_gives_plain_differences():
    sp = tp = np.zeros((1, 3))
    p = np.array([[1, 2], [3, 4]], np.float32)
    q = np.array([[1.5, 2], [3, 3]], np.float32)
    res, js, jt = rp.evaluate(0, sp, tp, [0, 2], p, q)
    assert np.array_equal(res, (q - p).astype(np.float64))


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [0, 1])
def test_gpu_vs_oracle(kind):
    from nautilus_b200 import CorrelativeScanMatcher
    from nautilus_b200.residuals import evaluate_residual_blocks
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    sp, tp, offs, p, q, ns, nt = make_blocks(300, seed=2 + kind)
    res, js, jt = evaluate_residual_blocks(m, kind, sp, tp, offs, p, q, ns, nt)
    wres, wjs, wjt = rp.evaluate(kind, sp, tp, offs, p, q, ns, nt)
    for got, want in ((res, wres), (js, wjs), (jt, wjt)):
        assert np.all(np.abs(got - want) <= TOL * np.maximum(1.0, np.abs(want)))
    only = evaluate_residual_blocks(m, kind, sp, tp, offs, p, q, ns, nt, jacobians=False)
    assert np.array_equal(only, res)
    # empty batch / block
    z = evaluate_residual_blocks(m, kind, np.zeros((1, 3)), np.zeros((1, 3)), [0, 0],
                                 np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 2)))
    assert z[0].shape == (0, 2)
    m.close()


@pytest.mark.gpu
def test_gpu_large_batch_linearity():
    """1e6 correspondences: at identical poses the point residual is target - R-invariant
    difference, so scaling both clouds scales the residuals (a size-independent property)."""
    from nautilus_b200 import CorrelativeScanMatcher
    from nautilus_b200.residuals import evaluate_residual_blocks
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    rng = np.random.default_rng(5)
    nb, per = 1000, 1000
    offs = np.arange(nb + 1, dtype=np.int64) * per
    poses = rng.uniform(-5, 5, size=(nb, 3))
    p = rng.uniform(-8, 8, size=(nb * per, 2)).astype(np.float32)
    q = rng.uniform(-8, 8, size=(nb * per, 2)).astype(np.float32)
    r1 = evaluate_residual_blocks(m, 0, poses, poses, offs, p, q, jacobians=False)
    r2 = evaluate_residual_blocks(m, 0, poses, poses, offs, 2 * p, 2 * q, jacobians=False)
    assert np.allclose(r2, 2 * r1, rtol=0, atol=1e-12)
    assert np.allclose(r1, (q.astype(np.float64) - p.astype(np.float64)), rtol=0, atol=1e-13 * 16)
    m.close()
