"""GPU parity tests of the front end (frontend.cu) through the C ABI against the
CPU restatement (oracle/frontend_port.py, itself pinned against the compiled
reference) and the reference's golden outputs.  Bar: points and counts bit-exact
(float products and compaction order are the reference's); scatter scores within
1e-6 relative (float sums in the reference's order, closed-form eigenvalues)."""
import os

import numpy as np
import pytest

from nautilus_b200 import CorrelativeScanMatcher, synth, synth_ranges as sr
from oracle import frontend_port as fp

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def matcher():
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    yield m
    m.close()


@pytest.mark.parametrize("tr", [0, 1])
def test_csm_variant_vs_reference_golden(matcher, tr):
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "frontend_golden.npz"))
    matcher.upload_ranges(g["ranges"], sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN,
                          sr.MAX_RANGE, variant=1, truncate_ends=bool(tr))
    xy, offs = matcher.download_scans()
    assert np.array_equal(offs, g[f"csm_t{tr}_offsets"])
    assert np.array_equal(bits(xy), bits(g[f"csm_t{tr}_xy"]))


def test_nautilus_variant_and_keyframes_vs_reference_golden(matcher):
    """nautilus::LaserScanToPointCloud (src/input/pointcloud_helpers.cc:28-48), the scatter
    score and GetLCCandidates (lc_candidate_filter.cc:34-81) against fixtures produced by
    the reference's own translation units: points bit-exact, candidates identical, scores
    to 1e-6 (float sums in the reference's order; eigenvalues in closed form)."""
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "frontend_golden.npz"))
    matcher.upload_ranges(g["ranges"], sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN,
                          sr.MAX_RANGE, variant=0)
    xy, offs = matcher.download_scans()
    assert np.array_equal(offs, g["nautilus_offsets"])
    assert np.array_equal(bits(xy), bits(g["nautilus_xy"]))
    matcher.upload_scans(g["lc_xy"], g["lc_offsets"])
    assert np.allclose(matcher.scatter_scores(), g["lc_scores"], rtol=1e-6, atol=0)
    cand = matcher.GetLCCandidates(g["lc_poses"][:, :2].astype(np.float32))
    assert cand.tolist() == g["lc_candidates"].tolist()


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("n_beams", [1080, 37, 1])
def test_vs_port_seeded(matcher, variant, n_beams):
    r = sr.make_ranges(64, n_beams=n_beams, seed=20 + n_beams)
    inc = np.float32(1.5 * np.pi / n_beams)
    amax = np.float32(sr.ANGLE_MIN + inc * np.float32(n_beams))
    matcher.upload_ranges(r, sr.ANGLE_MIN, amax, inc, sr.RANGE_MIN, sr.MAX_RANGE, variant=variant,
                          truncate_ends=True)
    xy, offs = matcher.download_scans()
    for s in range(len(r)):
        if variant == 0:
            want = fp.laser_to_points_nautilus(r[s], sr.ANGLE_MIN, inc, sr.RANGE_MIN, sr.MAX_RANGE)
        else:
            want = fp.laser_to_points_csm(r[s], sr.ANGLE_MIN, amax, inc, sr.RANGE_MIN, sr.MAX_RANGE, True)
        got = xy[offs[s]:offs[s + 1]]
        assert got.shape == want.shape, (s, got.shape, want.shape)
        assert np.array_equal(bits(got), bits(want)), s


def test_empty_batches(matcher):
    matcher.upload_ranges(np.zeros((0, 1080), np.float32), sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC,
                          sr.RANGE_MIN, sr.MAX_RANGE)
    xy, offs = matcher.download_scans()
    assert len(xy) == 0 and offs.tolist() == [0]
    matcher.upload_ranges(np.full((3, 16), 99.0, np.float32), sr.ANGLE_MIN, sr.ANGLE_MAX,
                          sr.ANGLE_INC, sr.RANGE_MIN, sr.MAX_RANGE, variant=1)
    xy, offs = matcher.download_scans()
    assert len(xy) == 0 and offs.tolist() == [0, 0, 0, 0]


def test_large_batch_roundtrip_properties(matcher):
    """4096 scans: counts equal the number of in-range beams, every point's norm is its
    range to float rounding, order follows the beam index (variant 0)."""
    r = sr.make_ranges(4096, seed=9)
    matcher.upload_ranges(r, sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN, sr.MAX_RANGE)
    xy, offs = matcher.download_scans()
    with np.errstate(invalid="ignore"):
        ok = (r >= sr.RANGE_MIN) & (r <= sr.MAX_RANGE)
    assert np.array_equal(offs[1:] - offs[:-1], ok.sum(axis=1))
    norm = np.hypot(xy[:, 0].astype(np.float64), xy[:, 1].astype(np.float64))
    assert np.allclose(norm, r[ok].astype(np.float64), rtol=3e-7, atol=0)


def test_matching_after_device_conversion_equals_host_upload(matcher):
    """Clouds produced on the device feed the matcher exactly like uploaded clouds."""
    r = sr.make_ranges(10, seed=31, dirty=True)
    # two views of one place: the second scan is the first, shifted by a few beams
    r[1, :-7] = r[0, 7:]
    matcher.upload_ranges(r, sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN, sr.MAX_RANGE,
                          variant=1)
    got = matcher.match_batch([0, 2], [1, 3], -0.3, 0.3, with_stats=True)
    xy, offs = matcher.download_scans()
    m2 = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m2.upload_scans(xy, offs)
    want = m2.match_batch([0, 2], [1, 3], -0.3, 0.3, with_stats=True)
    m2.close()
    for a, b in zip(got, want):
        assert np.array_equal(a, b)


def test_scatter_scores_and_candidates(matcher):
    xy, offs, poses = synth.make_scan_set(60, group=3, seed=4, return_poses=True)
    matcher.upload_scans(xy, offs)
    got = matcher.scatter_scores()
    want = np.array([fp.scatter_score(xy[offs[s]:offs[s + 1]]) for s in range(60)])
    assert np.allclose(got, want, rtol=1e-6, atol=0)
    # a trajectory: spread the poses out so that the distance rule matters
    traj = (poses[:, :2] + np.arange(60)[:, None] * np.array([1.3, 0.4])).astype(np.float32)
    cand = matcher.GetLCCandidates(traj)
    assert cand.tolist() == fp.lc_candidates(traj, want).tolist()
    assert len(cand) >= 1


def test_nn_index_after_device_conversion(matcher):
    """The NN path fetches device-made clouds for its host KD-tree layout."""
    r = sr.make_ranges(6, seed=8, dirty=False)
    matcher.upload_ranges(r, sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN, sr.MAX_RANGE)
    matcher.nn_build(0.25)
    ident = np.tile(np.array([1, 0, 0, 0, 1, 0], np.float32), (2, 1))
    idx, offs = matcher.nn_match_batch([0, 1], [0, 1], ident, 0.25)[:2]
    # a cloud matched against itself with the identity: every point finds itself
    n0 = offs[1] - offs[0]
    assert np.array_equal(idx[:n0], np.arange(n0))
