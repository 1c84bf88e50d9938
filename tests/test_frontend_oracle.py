"""CPU tests of the front-end oracle (oracle/frontend_port.py): the numpy
restatement against golden outputs of the compiled reference
(tests/golden/frontend_golden.npz) and, when oracle/_ref is present, against the
reference itself; plus the host keyframe selection of the C-ABI library."""
import os

import numpy as np
import pytest

from nautilus_b200 import synth_ranges as sr
from oracle import frontend_port as fp


@pytest.fixture(scope="module")
def fe_golden():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "frontend_golden.npz"))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("tr", [0, 1])
def test_port_vs_reference_golden(fe_golden, tr):
    r = fe_golden["ranges"]
    offs = fe_golden[f"csm_t{tr}_offsets"]
    xy = fe_golden[f"csm_t{tr}_xy"]
    for s in range(len(r)):
        got = fp.laser_to_points_csm(r[s], sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN,
                                     sr.MAX_RANGE, bool(tr))
        want = xy[offs[s]:offs[s + 1]]
        assert got.shape == want.shape
        assert np.array_equal(bits(got), bits(want))
    assert offs[4] - offs[3] == 0          # the scan without a valid return is empty


def test_port_nautilus_variant_and_scatter_vs_reference_golden(fe_golden):
    """The variant the Nautilus node calls (src/input/pointcloud_helpers.cc:28-48) and
    the keyframe filter (lc_candidate_filter.cc:34-81): port against fixtures made by the
    reference's own translation units."""
    g = fe_golden
    r, offs, xy = g["ranges"], g["nautilus_offsets"], g["nautilus_xy"]
    for s in range(len(r)):
        got = fp.laser_to_points_nautilus(r[s], sr.ANGLE_MIN, sr.ANGLE_INC, sr.RANGE_MIN, sr.MAX_RANGE)
        want = xy[offs[s]:offs[s + 1]]
        assert got.shape == want.shape and np.array_equal(bits(got), bits(want)), s
    lo = g["lc_offsets"]
    scores = np.asarray([fp.scatter_score(g["lc_xy"][lo[k]:lo[k + 1]]) for k in range(len(lo) - 1)])
    assert np.array_equal(scores, g["lc_scores"])
    cand = fp.lc_candidates(g["lc_poses"][:, :2].astype(np.float32), scores)
    assert cand.tolist() == g["lc_candidates"].tolist() and len(cand) >= 3


def test_port_vs_compiled_reference(ref_lib):
    r = sr.make_ranges(40, seed=11)
    for s in range(len(r)):
        for tr in (False, True):
            want = ref_lib.csm_laser_to_points(r[s], sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC,
                                               sr.RANGE_MIN, sr.MAX_RANGE, tr)
            got = fp.laser_to_points_csm(r[s], sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC,
                                         sr.RANGE_MIN, sr.MAX_RANGE, tr)
            assert got.shape == want.shape and np.array_equal(bits(got), bits(want))
        want = ref_lib.nautilus_laser_to_points(r[s], sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC,
                                                sr.RANGE_MIN, sr.MAX_RANGE)
        got = fp.laser_to_points_nautilus(r[s], sr.ANGLE_MIN, sr.ANGLE_INC, sr.RANGE_MIN, sr.MAX_RANGE)
        assert got.shape == want.shape and np.array_equal(bits(got), bits(want))
        if len(want) > 1:
            assert fp.scatter_score(want) == ref_lib.scatter_score(want)


def test_nautilus_variant_shares_the_arithmetic():
    """Without dropped beams and with an increment whose float multiples equal its
    float running sums (a power of two), the two variants see the same angles and
    differ only by the csm angle-window test and glancing filter: the nautilus
    variant's points contain the csm variant's, bit for bit, in order."""
    n = 512
    inc = np.float32(2.0 ** -8)
    amin = np.float32(-1.0)
    amax = np.float32(amin + inc * np.float32(n))
    r = sr.make_ranges(3, n_beams=n, seed=2, dirty=False)
    for s in range(len(r)):
        a = fp.laser_to_points_nautilus(r[s], amin, inc, 0.05, 30.0)
        b = fp.laser_to_points_csm(r[s], amin, amax, inc, 0.05, 30.0, False)
        assert len(a) == n
        ia = {tuple(v) for v in bits(a).tolist()}
        assert all(tuple(v) in ia for v in bits(b).tolist())


def test_scatter_score_and_candidates():
    rng = np.random.default_rng(0)
    line = np.stack([np.linspace(-5, 5, 400), 0.01 * rng.normal(size=400)], 1).astype(np.float32)
    blob = rng.normal(size=(400, 2)).astype(np.float32)
    assert fp.scatter_score(line) < 0.01
    assert 0.7 < fp.scatter_score(blob) <= 1.0
    scores = np.array([0.9, 0.9, 0.2, 0.9, 0.9])
    poses = np.array([[0, 0], [1, 0], [6, 0], [7, 0], [12.5, 0]], np.float32)
    # 1 is too close to 0; 2 is far enough but scores low; 3 is 7 m from 0; 4 is 5.5 m from 3
    assert fp.lc_candidates(poses, scores).tolist() == [0, 3, 4]


def test_library_keyframe_selection_matches_port():
    """nautilus_lc_candidates is host code of the C-ABI library: no GPU needed."""
    import ctypes as C
    from nautilus_b200 import _lib
    L = _lib.lib()
    rng = np.random.default_rng(3)
    n = 500
    poses = np.cumsum(rng.normal(0, 0.6, size=(n, 2)), axis=0).astype(np.float32)
    scores = rng.uniform(0.4, 1.0, size=n)
    out = np.empty(n, np.int64)
    m = C.c_int32()
    assert L.nautilus_lc_candidates(_lib.ptr(scores), _lib.ptr(poses), n, 5.0, 0.70, _lib.ptr(out),
                                    C.byref(m)) == 0
    assert out[:m.value].tolist() == fp.lc_candidates(poses, scores).tolist()
    assert m.value > 3
