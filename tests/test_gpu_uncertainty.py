"""GPU parity of GetUncertaintyMatrix (CorrelativeScanMatcher.cc:267-373) against
outputs of the compiled reference (tests/golden/unc_golden.npz) and, when
present, the compiled reference itself.  Tolerance 1e-4 relative to the largest
matrix entry (float moment sums of exp(cost); CUDA exp/log
differ from glibc's by ulps), written here."""
import os

import numpy as np
import pytest

from nautilus_b200 import CorrelativeScanMatcher, CsmError, synth

pytestmark = pytest.mark.gpu
TOL = 1e-4


@pytest.fixture(scope="module")
def unc():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "unc_golden.npz"))


def close(got, want):
    return float(np.abs(got - want).max() / np.abs(want).max())


def test_uncertainty_2x2_vs_golden(unc):
    m = CorrelativeScanMatcher(*unc["params"])
    for k in range(3):
        got = m.GetUncertaintyMatrix(unc[f"a_{k}"], unc[f"b_{k}"], float(unc["rot"][k]))
        assert got.shape == (2, 2) and got.dtype == np.float32
        assert close(got, unc["u2"][k]) <= TOL, (k, got, unc["u2"][k])
    m.close()


def test_uncertainty_3x3_vs_golden(unc):
    m = CorrelativeScanMatcher(*unc["params"])
    for k in range(2):
        got = m.GetUncertaintyMatrix(unc[f"a_{k}"], unc[f"b_{k}"])
        assert got.shape == (3, 3)
        assert close(got, unc["u3"][k]) <= TOL, (k, got, unc["u3"][k])
    m.close()


def test_uncertainty_vs_compiled_reference(ref_lib):
    a, b, truth = synth.make_pair(77, n_beams=540)
    rm = ref_lib.RefMatcher(30, 2, 0.3, 0.03)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    for rot in (0.0, float(np.float32(truth[2])), -0.3):
        assert close(m.GetUncertaintyMatrix(a, b, rot), rm.uncertainty2(a, b, rot)) <= TOL
    # GetLocalUncertaintyStats (cc:375-398) from the reference's own pieces
    clouds = [b, synth.make_pair(78, n_beams=540)[1]]
    cond, scale = m.GetLocalUncertaintyStats(clouds, a)
    want_c, want_s = 0.0, 0.0
    for c in clouds:
        r = rm.match(a, c)
        u = rm.uncertainty2(a, c, float(r[3])).astype(np.float64)
        ev = np.linalg.eigvalsh((u + u.T) / 2)
        want_c += ev.max() / ev.min()
        want_s += ev.max()
    assert abs(cond - want_c / 2) <= 1e-3 * abs(want_c / 2)
    assert abs(scale - want_s / 2) <= 1e-3 * abs(want_s / 2)
    m.close()


def test_uncertainty_errors():
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    with pytest.raises(CsmError):
        m.GetUncertaintyMatrix(np.zeros((0, 2), np.float32), np.ones((3, 2), np.float32), 0.0)
    m.close()


def test_uncertainty_batch_equals_single(unc):
    """nautilus_csm_uncertainty_batch over uploaded scans = the single-pair call, pair by
    pair (same kernels' arithmetic, tables built once per target), incl. several table waves."""
    xy, offs = synth.make_scan_set(12, group=4, seed=8, n_beams=360)
    pa, pb = synth.make_pair_list(12, group=4)
    rot = np.linspace(-0.4, 0.4, len(pa))
    for max_tables in (0, 2):
        m = CorrelativeScanMatcher(30, 2, 0.3, 0.03, max_tables=max_tables)
        m.upload_scans(xy, offs)
        got = m.uncertainty_batch(pa, pb, rot)
        for k in range(0, len(pa), 5):
            a, b = xy[offs[pa[k]]:offs[pa[k] + 1]], xy[offs[pb[k]]:offs[pb[k] + 1]]
            assert np.array_equal(got[k], m.GetUncertaintyMatrix(a, b, float(rot[k]))), k
        m.close()
