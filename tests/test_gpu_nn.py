"""GPU parity tests of the point-to-point nearest-neighbour path (K7), called
through the C ABI, against outputs of the compiled reference KD-tree
(tests/golden/nn_golden.npz), the CPU restatement (oracle/nn_oracle.cc) and,
when present, the compiled reference itself (oracle/_ref).

Bar: matched target indices bit-exact (index work), returned distances
bit-exact where matched (they are the reference's float distances)."""
import numpy as np
import pytest

from nautilus_b200 import (CorrelativeScanMatcher, CsmError, GetPointToPointMatching,
                           pose_transform, synth)
from oracle import nn_port

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nn_golden():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "nn_golden.npz"))


def make_matcher(xy, offs, cell):
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m.upload_scans(xy, offs)
    m.nn_build(cell)
    return m


@pytest.mark.parametrize("name", ["scan", "adv"])
def test_matches_vs_reference_golden(nn_golden, name):
    g = nn_golden
    m = make_matcher(g[f"{name}_xy"], g[f"{name}_offsets"], 0.7)
    walked = 0
    for k, thr in enumerate(g["thresholds"]):
        idx, offs, dist = m.nn_match_batch(g[f"{name}_src"], g[f"{name}_tgt"], g[f"{name}_m"],
                                           float(thr), with_dist=True)
        want = g[f"{name}_idx_{k}"]
        assert np.array_equal(offs, g[f"{name}_out_offsets"])
        assert np.array_equal(idx, want), (name, thr, int((idx != want).sum()))
        hit = want >= 0
        assert np.array_equal(dist[hit], g[f"{name}_dist_{k}"][hit])
        walked += m.nn_last_stats()["n_walked"]
    if name == "adv":
        assert walked > 0  # lattice / duplicate clouds must exercise the exact tree walk
    assert np.array_equal(m.nn_self_index(), g[f"{name}_self_idx"])
    m.close()


def test_pose_transform_matches_golden_and_float64(nn_golden):
    g = nn_golden
    for p, (s, t) in enumerate(zip(g["scan_src"][:40], g["scan_tgt"][:40])):
        got = pose_transform(g["scan_poses"][s], g["scan_poses"][t])
        assert np.array_equal(got, g["scan_m"][p])
        (xs, ys, ts), (xt, yt, tt) = g["scan_poses"][s], g["scan_poses"][t]
        c, sn = np.cos(ts - tt), np.sin(ts - tt)
        ct, st = np.cos(tt), np.sin(tt)
        dx, dy = xs - xt, ys - yt
        want = [c, -sn, ct * dx + st * dy, sn, c, -st * dx + ct * dy]
        assert np.allclose(got, want, atol=2e-6)


def test_random_pairs_vs_port_and_quirk():
    """Fresh seeds at BASELINE beam count; the tree's single-child corner must
    show up (reference != exact nearest neighbour) and be reproduced."""
    xy, offs, poses = synth.make_scan_set(40, group=5, seed=3, return_poses=True)
    rng = np.random.default_rng(9)
    est = poses + np.concatenate([rng.normal(0, 0.1, (40, 2)), rng.normal(0, 0.03, (40, 1))], 1)
    src, tgt = synth.make_pair_list(40, group=5)
    tr = np.stack([pose_transform(est[s], est[t]) for s, t in zip(src, tgt)])
    m = make_matcher(xy, offs, 0.25)
    idx, o, dist = m.nn_match_batch(src, tgt, tr, 0.25, with_dist=True)
    want, wo, wdist = nn_port.match_batch(xy, offs, src, tgt, tr, 0.25)
    assert np.array_equal(o, wo) and np.array_equal(idx, want)
    assert np.array_equal(dist[want >= 0], wdist[want >= 0])
    exact = nn_port.brute_force(xy, offs, src, tgt, tr, 0.25)
    assert (exact != want).any()
    stats = m.nn_last_stats()
    assert stats["n_walked"] < 0.01 * len(idx)  # the grid answers all but the ambiguous queries
    m.close()


def test_against_compiled_reference(ref_lib):
    xy, offs, poses = synth.make_scan_set(30, group=5, seed=77, n_beams=720, return_poses=True)
    src, tgt = synth.make_pair_list(30, group=5)
    tr = np.stack([pose_transform(poses[s], poses[t]) for s, t in zip(src, tgt)])
    m = make_matcher(xy, offs, 1.0)
    for thr in (1.0, 0.25, 0.1):
        idx, o = m.nn_match_batch(src, tgt, tr, thr)
        _, want, wo = ref_lib.nn_match_batch(xy, offs, src, tgt, tr, thr)
        assert np.array_equal(o, wo) and np.array_equal(idx, want), thr
    m.close()


def test_get_point_to_point_matching_struct():
    xy, offs, poses = synth.make_scan_set(5, group=5, seed=5, n_beams=360, return_poses=True)
    rng = np.random.default_rng(1)
    normals = rng.normal(size=xy.shape).astype(np.float32)
    m = make_matcher(xy, offs, 0.25)
    c = GetPointToPointMatching(m, xy, offs, normals, poses, 1, 3, 0.25)
    want, wo, _ = nn_port.match_batch(xy, offs, [1], [3], pose_transform(poses[1], poses[3]), 0.25)
    hit = np.nonzero(want >= 0)[0]
    assert (c.source_index, c.target_index) == (1, 3)
    assert np.array_equal(c.source_points, xy[offs[1] + hit])
    assert np.array_equal(c.target_points, xy[offs[3] + want[hit]])
    assert np.array_equal(c.source_normals, normals[offs[1] + hit])   # no duplicates: self index
    assert np.array_equal(c.target_normals, normals[offs[3] + want[hit]])
    d = c.source_points - c.target_points
    ref_diff = 0.0
    for v in np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(np.float32)):
        ref_diff += float(v)
    assert c.difference == ref_diff
    m.close()


def test_nn_errors_and_edges():
    xy = np.asarray([[0, 0], [1, 0], [5, 5]], np.float32)
    offs = np.asarray([0, 2, 2, 3], np.int64)   # scan 1 is empty
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m.upload_scans(xy, offs)
    ident = np.asarray([1, 0, 0, 0, 1, 0], np.float32)
    with pytest.raises(CsmError):     # build first
        m.nn_match_batch([0], [0], ident, 0.25)
    m.nn_build(0.25)
    with pytest.raises(CsmError):     # threshold above the grid cell
        m.nn_match_batch([0], [0], ident, 0.5)
    with pytest.raises(CsmError):
        m.nn_match_batch([0], [7], ident, 0.25)
    idx, o = m.nn_match_batch([0, 1, 0, 2], [0, 0, 1, 0], np.tile(ident, (4, 1)), 0.25)
    assert o.tolist() == [0, 2, 2, 4, 5]
    assert idx.tolist() == [0, 1, -1, -1, -1]   # self, empty source, empty target, too far
    idx, o = m.nn_match_batch([], [], np.zeros((0, 6), np.float32), 0.25)
    assert len(idx) == 0 and o.tolist() == [0]
    m.upload_scans(xy, offs)          # a new scan set invalidates the index
    with pytest.raises(CsmError):
        m.nn_match_batch([0], [0], ident, 0.25)
    m.close()


@pytest.mark.parametrize("kind", [0, 1])
def test_resident_chain_equals_host_chain(kind):
    """nn_match_resident -> evaluate_resident_blocks (correspondences never leave the
    device) against the host-buffer chain nn_match_batch -> AddCorrespondence gather ->
    evaluate_residual_blocks, which is itself pinned to the oracles above: same
    correspondences in the same order, residuals and Jacobians bit for bit.  Duplicate
    points make the normal lookup go through GetPointNormal's self index."""
    from nautilus_b200.point_matching import correspondences_from_matches
    from nautilus_b200.residuals import evaluate_residual_blocks, evaluate_resident_blocks
    xy, offs, poses = synth.make_scan_set(12, group=4, seed=9, n_beams=360, return_poses=True)
    xy = xy.copy()
    for s in range(12):               # a few exact duplicates per scan
        o = int(offs[s])
        if offs[s + 1] - o > 40:
            xy[o + 30] = xy[o + 10]
            xy[o + 31] = xy[o + 10]
    rng = np.random.default_rng(3)
    normals = rng.normal(size=xy.shape).astype(np.float32)
    m = make_matcher(xy, offs, 0.25)
    m.nn_upload_normals(normals)
    src = np.asarray([0, 1, 2, 3, 5, 4, 8, 9, 10, 11, 0, 7], np.int32)
    tgt = np.asarray([1, 0, 3, 2, 4, 5, 9, 8, 11, 10, 6, 7], np.int32)
    tf = np.stack([pose_transform(poses[s], poses[t]) for s, t in zip(src, tgt)])
    tf[10] = [1, 0, 1000, 0, 1, 1000]   # a pair without a single match: an empty block
    # host chain
    idx, out_offs = m.nn_match_batch(src, tgt, tf, 0.25)
    cs = correspondences_from_matches(m, xy, offs, normals, poses, src, tgt, idx, out_offs)
    counts = np.asarray([len(c.source_points) for c in cs], np.int64)
    want_offs = np.concatenate([[0], np.cumsum(counts)])
    cat = lambda k: np.concatenate([getattr(c, k) for c in cs])  # noqa: E731
    sp, tp = poses[src] + rng.normal(scale=0.05, size=(len(src), 3)), poses[tgt]
    want = evaluate_residual_blocks(m, kind, sp, tp, want_offs, cat("source_points"),
                                    cat("target_points"), cat("source_normals"), cat("target_normals"))
    # resident chain
    corr_offs = m.nn_match_resident(src, tgt, tf, 0.25)
    assert np.array_equal(corr_offs, want_offs)
    assert corr_offs[-1] > 1000 and counts[10] == 0
    got = evaluate_resident_blocks(m, kind, sp, tp, corr_offs)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    # page-locked result buffers (nautilus_host_alloc), reused by a second call
    pin = evaluate_resident_blocks(m, kind, sp, tp, corr_offs, pinned=True)
    pin = evaluate_resident_blocks(m, kind, sp, tp, corr_offs, out=pin)
    for g, w in zip(pin, want):
        assert np.array_equal(g, w)
    del pin
    # re-evaluation with moved poses, residuals only (what a Ceres iteration does)
    sp2 = sp + 0.01
    r2 = evaluate_resident_blocks(m, kind, sp2, tp, corr_offs, jacobians=False)
    w2 = evaluate_residual_blocks(m, kind, sp2, tp, want_offs, cat("source_points"), cat("target_points"),
                                  cat("source_normals"), cat("target_normals"), jacobians=False)
    assert np.array_equal(r2, w2)
    m.close()


def test_resident_chain_errors():
    xy, offs = synth.make_scan_set(4, group=4, seed=2, n_beams=120)
    m = make_matcher(xy, offs, 0.25)
    from nautilus_b200.residuals import evaluate_resident_blocks
    ident = np.asarray([1, 0, 0, 0, 1, 0], np.float32)
    with pytest.raises(CsmError):     # nothing resident yet
        evaluate_resident_blocks(m, 0, np.zeros((1, 3)), np.zeros((1, 3)), np.asarray([0, 1]))
    co = m.nn_match_resident([0], [1], ident, 0.25)
    with pytest.raises(CsmError):     # normal residuals need the normals
        evaluate_resident_blocks(m, 1, np.zeros((1, 3)), np.zeros((1, 3)), co)
    r = evaluate_resident_blocks(m, 0, np.zeros((1, 3)), np.zeros((1, 3)), co, jacobians=False)
    assert r.shape == (int(co[-1]), 2)
    co = m.nn_match_resident([], [], np.zeros((0, 6), np.float32), 0.25)
    assert co.tolist() == [0]
    m.close()


def test_planar_subsets_through_resident_chain():
    """PointcloudType::PLANAR / EDGE (solver.cc:96-107): source points and the lookup tree
    come from the planar subsets (<= 20 points per scan), the normals from the FULL clouds'
    trees (solver.cc:152-155).  The subsets are uploaded as a scan set of their own; their
    normals are gathered once through the full clouds' self index (GetPointNormal depends on
    the coordinates only).  Checked against the oracle's search on the subsets."""
    from nautilus_b200.residuals import evaluate_residual_blocks, evaluate_resident_blocks
    xy, offs, poses = synth.make_scan_set(8, group=4, seed=13, n_beams=360, return_poses=True)
    rng = np.random.default_rng(7)
    normals = rng.normal(size=xy.shape).astype(np.float32)
    full = make_matcher(xy, offs, 0.25)
    self_full = full.nn_self_index()
    # planar subsets: 20 evenly spaced points of every scan
    pick = [int(offs[s]) + np.linspace(0, offs[s + 1] - offs[s] - 1, 20).astype(np.int64) for s in range(8)]
    sub_xy = np.concatenate([xy[p] for p in pick])
    sub_offs = np.arange(9, dtype=np.int64) * 20
    sub_normals = np.concatenate([normals[int(offs[s]) + self_full[pick[s]]] for s in range(8)])
    m = make_matcher(sub_xy, sub_offs, 0.25)
    m.nn_upload_normals(sub_normals)
    src = np.asarray([0, 1, 2, 3, 4, 5, 6, 7], np.int32)
    tgt = np.asarray([1, 0, 3, 2, 5, 4, 7, 6], np.int32)
    tf = np.stack([pose_transform(poses[s], poses[t]) for s, t in zip(src, tgt)])
    co = m.nn_match_resident(src, tgt, tf, 0.25)
    want_idx, wo, _ = nn_port.match_batch(sub_xy, sub_offs, src, tgt, tf, 0.25)
    hit = want_idx >= 0
    blk = np.searchsorted(wo, np.arange(len(want_idx)), side="right") - 1
    assert np.array_equal(co, np.concatenate([[0], np.cumsum(np.bincount(blk[hit], minlength=8))]))
    assert co[-1] > 0
    sp = sub_xy[sub_offs[src[blk[hit]]] + (np.arange(len(want_idx)) - wo[blk])[hit]]
    tp = sub_xy[sub_offs[tgt[blk[hit]]] + want_idx[hit]]
    sn = sub_normals[sub_offs[src[blk[hit]]] + (np.arange(len(want_idx)) - wo[blk])[hit]]
    tn = sub_normals[sub_offs[tgt[blk[hit]]] + want_idx[hit]]
    got = evaluate_resident_blocks(m, 1, poses[src], poses[tgt], co)
    want = evaluate_residual_blocks(m, 1, poses[src], poses[tgt], co, sp, tp, sn, tn)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    m.close()
    full.close()


def test_async_index_build_overlapping_a_csm_batch():
    """nautilus_nn_build_begin runs the host KD-tree layout on library threads while the
    caller matches a CSM batch; the joined index gives the matches a plain build gives,
    and replacing the scan set while a build is in flight is safe."""
    xy, offs, poses = synth.make_scan_set(40, group=5, seed=13, n_beams=360, return_poses=True)
    pa, pb = synth.make_pair_list(40, group=5)
    tr = np.stack([pose_transform(poses[a], poses[b]) for a, b in zip(pa, pb)])
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m.upload_scans(xy, offs)
    m.nn_build(0.25)
    want_idx, want_off = m.nn_match_batch(pa, pb, tr, 0.25)
    want_self = m.nn_self_index()
    m.upload_scans(xy, offs)          # drops the index
    m.nn_build_begin(0.25)
    csm = m.match_batch(pa, pb, -0.5, 0.5)   # overlaps the host build
    m.nn_build(0.25)                  # joins
    idx, off = m.nn_match_batch(pa, pb, tr, 0.25)
    assert np.array_equal(idx, want_idx) and np.array_equal(off, want_off)
    assert np.array_equal(m.nn_self_index(), want_self)
    m.nn_build_begin(0.25)
    m.upload_scans(xy[:offs[10]], offs[:11])  # cancels the build in flight
    m.nn_build(0.5)
    assert len(m.nn_self_index()) == offs[10]
    assert len(csm[0]) == len(pa)
    m.close()
