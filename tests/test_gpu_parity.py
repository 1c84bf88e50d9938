"""GPU parity tests: the sm_100a path, called through the C ABI, against the
oracle port and the committed reference fixtures.

Bars (north_star): argmax pose bit-exact (float x, float y, float theta); score
and table values within 1e-5 relative (tables: relative to max(|ref|, 1e-10),
the lookup floor of CorrelativeScanMatcher.h:20).  With margin_cells=-1 the
tables are additionally required to be bit-identical."""
import numpy as np
import pytest

from nautilus_b200 import CorrelativeScanMatcher, CsmError, LookupTable, synth
from oracle import port

pytestmark = pytest.mark.gpu

REL = 1e-5


def table_close(got, want):
    err = np.abs(got - want) / np.maximum(np.abs(want), 1e-10)
    return float(err.max())


def pose(r):
    return (r[1][0][0], r[1][0][1], r[1][1])


# ---- the reference's unit tests through the GPU library ----------------------
# (third_party/csm/test/correlative_scan_matcher_tests.cc:12-106)
def test_ref_unit_cell_mapping():
    t = LookupTable(4, 1)
    t.SetPointValue((0, 0), 0.5)
    assert t.GetPointValue((0, 0)) == 0.5
    assert t.GetPointValue((.5, .5)) == 0.5
    assert t.GetPointValue((-.5, -.5)) < 0.5
    t = LookupTable(4, 1)
    t.SetPointValue((-3.9, -3.9), 0.01)
    t.SetPointValue((-3.0, -3.0), 0.275)
    assert t.GetPointValue((-3.51, -3.51)) == 0.01
    assert t.GetPointValue((-3.49, -3.49)) == 0.01
    assert t.GetPointValue((-2.99, -2.99)) == 0.275
    t = LookupTable(4, 1)
    t.SetPointValue((-4, -4), 0.23)
    assert t.GetPointValue((-4, -3.99)) == 0.23
    t.SetPointValue((4, -4), 0.24)
    assert t.GetPointValue((3.99, -4)) < 0.24


def test_ref_unit_low_res_simple():
    t = LookupTable(4, 1)
    t.SetPointValue((0, 0), 1)
    m = CorrelativeScanMatcher(4, 4, 1, 1)
    assert m.GetLookupTableLowRes(t).GetPointValue((0, 0)) == 1


def test_ref_unit_low_res():
    t = LookupTable(4, 0.5)
    t.SetPointValue((0, 0.1), 0.5)
    t.SetPointValue((0, 0.7), 0.56)
    m = CorrelativeScanMatcher(4, 4, 1, 0.5)
    assert m.GetLookupTableLowRes(t).GetPointValue((0, 0)) == 0.56


def test_ref_unit_low_res_inexpressible():
    t = LookupTable(0.3, 0.03)
    t.SetPointValue((0, 0), 0.5)
    t.SetPointValue((0, 0.031), 0.56)
    m = CorrelativeScanMatcher(0.3, 0.3, 0.3, 0.03)
    low = m.GetLookupTableLowRes(t)
    assert (low.width, low.height) == (2, 2)
    assert low.GetPointValue((0, 0)) == 0.56
    t = LookupTable(0.3, 0.03)
    t.SetPointValue((0, 0), 0.5)
    t.SetPointValue((0, -0.031), 0.56)
    t.SetPointValue((-0.061, -0.031), 0.56)
    t.SetPointValue((-0.031, -0.031), 0.75)
    low = m.GetLookupTableLowRes(t)
    assert low.GetPointValue((0, 0)) == 0.5
    assert low.GetPointValue((-0.03, -0.03)) == 0.75
    assert low.GetPointValue((-0.03, 0.03)) == 1e-10
    assert low.GetPointValue((0.0, -0.03)) == 0.56


@pytest.mark.parametrize("cfg", [(30, 2, 0.3, 0.03), (30, 10, 0.3, 0.03), (4, 0.45, 0.3, 0.03),
                                 (30, 2, 0.5, 0.05), (10, 1, 0.2, 0.01)])
def test_regular_cell_rule_matches_exact_cells(cfg):
    """K5 / K6 take the cells of consecutive translations from the first one when its
    float quotient lies clear of a cell edge; the rule must agree with the exact map
    (floor((double)v / res), h:98-101) on random and edge-hugging coordinates."""
    m = CorrelativeScanMatcher(*cfg)
    n, bad_low, bad_high = m.regular_check(300000, seed=7)
    m.close()
    assert n > 1000000 and bad_low == 0 and bad_high == 0


def test_low_res_from_random_tables_vs_oracle():
    rng = np.random.default_rng(3)
    for (rg, lo, hi) in [(4, 1, 0.5), (0.3, 0.3, 0.03), (6, 0.3, 0.03), (5, 0.7, 0.1)]:
        m = CorrelativeScanMatcher(rg, 1, lo, hi)
        pm = port.PortMatcher(rg, 1, lo, hi)
        t = LookupTable(rg, hi, rng.uniform(-0.2, 1.0, size=(int(2 * rg / hi),) * 2))
        pt = port.PortTable(rg, hi, t.values.copy())
        assert np.array_equal(m.GetLookupTableLowRes(t).values, pm.lowres(pt).values)


# ---- tables -------------------------------------------------------------------
@pytest.mark.parametrize("margin", [-1, 0, 40])
def test_small_tables_vs_golden(golden, margin):
    m = CorrelativeScanMatcher(*golden["small_params"], margin_cells=margin)
    for k in range(4):
        hi, lo = m.GetLookupTables(golden[f"small_cloud_{k}"])
        if margin < 0:
            assert np.array_equal(hi.values, golden[f"small_hi_{k}"]), k
            assert np.array_equal(lo.values, golden[f"small_lo_{k}"]), k
        else:
            tol = REL if margin == 0 else 1e-3  # 40 cells: tails cut at 0.716^40
            assert table_close(hi.values, golden[f"small_hi_{k}"]) <= tol
            assert table_close(lo.values, golden[f"small_lo_{k}"]) <= tol


def test_full_size_tables_vs_oracle():
    pm = port.PortMatcher(30, 2, 0.3, 0.03)
    _, b, _ = synth.make_pair(21)
    ph = pm.highres(b)
    pl = pm.lowres(ph)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03, margin_cells=-1)
    hi, lo = m.GetLookupTables(b)
    assert np.array_equal(hi.values, ph.values) and np.array_equal(lo.values, pl.values)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    hi, lo = m.GetLookupTables(b)
    assert table_close(hi.values, ph.values) <= REL
    assert table_close(lo.values, pl.values) <= REL
    assert hi.values.max() == 1.0 and hi.values.min() >= 0.0


def test_tables_edge_inputs():
    pm = port.PortMatcher(4, 1, 0.5, 0.05)
    for margin in (-1, 0):
        m = CorrelativeScanMatcher(4, 1, 0.5, 0.05, margin_cells=margin)
        # empty cloud: all zero (normalize's m == M branch)
        hi, lo = m.GetLookupTables(np.zeros((0, 2), np.float32))
        assert not hi.values.any() and not lo.values.any()
        # every point outside the grid
        hi, lo = m.GetLookupTables(np.asarray([[9, 9], [-4.01, 0]], np.float32))
        assert not hi.values.any()
        # single point, corner points (Neumann seed + Triggs boundary), duplicates
        # non-finite coordinates fall outside the grid (NaN like the reference's
        # double -> uint64 conversion; +-inf / |v| >= 2^63: the reference's conversion is
        # undefined behaviour there, the library and the oracle drop the point)
        for pts in ([[0.0, 0.0]], [[-4.0, -4.0], [3.99, 3.99]], [[1, 1], [1, 1], [1.001, 1.001]],
                    [[-4.0, 3.99], [3.99, -4.0], [0, 0]],
                    [[np.nan, 0.5], [1.0, np.nan], [np.inf, 1.0], [1.0, -np.inf], [0.5, 0.25]]):
            pts = np.asarray(pts, np.float32)
            hi, lo = m.GetLookupTables(pts)
            ph = pm.highres(pts)
            if margin < 0:
                assert np.array_equal(hi.values, ph.values), pts
                assert np.array_equal(lo.values, pm.lowres(ph).values), pts
            else:
                assert table_close(hi.values, ph.values) <= REL


# ---- matching -----------------------------------------------------------------
@pytest.mark.parametrize("margin", [-1, 0])
def test_small_matches_vs_golden(golden, margin):
    m = CorrelativeScanMatcher(*golden["small_params"], margin_cells=margin)
    for k, row in enumerate(golden["small_match"]):
        a, b = golden[f"small_a_{k}"], golden[f"small_b_{k}"]
        for r, want in ((m.GetTransformation(a, b, -0.5, 0.5), row[:4]),
                        (m.GetTransformation(a, b), row[4:])):
            assert np.array_equal(np.asarray(pose(r), np.float64), want[1:]), (k, r, want)
            assert abs(r[0] - want[0]) <= REL * abs(want[0])


def test_std_matches_vs_golden(golden):
    m = CorrelativeScanMatcher(*golden["std_params"])
    for k, row in enumerate(golden["std_match"]):
        r = m.GetTransformation(golden[f"std_a_{k}"], golden[f"std_b_{k}"], -0.5, 0.5)
        assert np.array_equal(np.asarray(pose(r), np.float64), row[1:]), (k, r, row)
        assert abs(r[0] - row[0]) <= REL * abs(row[0])


def test_restricted_overload_vs_golden(golden):
    m = CorrelativeScanMatcher(*golden["small_params"])
    a, b = golden["small_a_0"], golden["small_b_0"]
    for ra, rb, rr, s, x, y, th in golden["small_restricted"]:
        r = m.GetTransformationRestricted(a, b, ra, rb, rr)
        assert np.array_equal(np.asarray(pose(r), np.float64), [x, y, th])
        assert abs(r[0] - s) <= REL * abs(s)


def test_full_size_matches_vs_oracle_with_counts():
    pm = port.PortMatcher(30, 2, 0.3, 0.03)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    for seed in range(30, 36):
        a, b, _ = synth.make_pair(seed)
        got, st = m.GetTransformation(a, b, -0.5, 0.5, with_stats=True)
        want, wst = pm.match(a, b, -0.5, 0.5, with_stats=True)
        assert pose(got) == (want[1], want[2], want[3]), (seed, got, want)
        assert abs(got[0] - want[0]) <= REL * abs(want[0])
        assert tuple(st) == tuple(int(v) for v in wst)


def test_match_edge_inputs():
    pm = port.PortMatcher(4, 1, 0.5, 0.05)
    m = CorrelativeScanMatcher(4, 1, 0.5, 0.05)
    a = np.asarray([[0.5, 0.5], [1.0, -1.0], [-2.0, 0.3]], np.float32)
    empty = np.zeros((0, 2), np.float32)
    far = np.asarray([[10.0, 0.0], [0.0, -9.0], [0.2, 0.1]], np.float32)
    a_nan = np.asarray([[0.5, 0.5], [np.nan, 1.0], [1.0, -1.0], [np.inf, 0.0], [-2.0, 0.3]], np.float32)
    # empty target: every candidate scores log(1e-10) -> all ties, all windows popped
    for qa, b, win in ((a, empty, (-0.1, 0.1)), (a, far, (-0.1, 0.1)), (a, a, (-0.2, 0.2)),
                       (a, a, (6.0, 6.5)), (a_nan, a, (-0.2, 0.2)), (a, a_nan, (-0.2, 0.2))):
        got, st = m.GetTransformation(qa, b, *win, with_stats=True)
        want, wst = pm.match(qa, b, *win, with_stats=True)
        assert pose(got) == (want[1], want[2], want[3]), (got, want)
        assert abs(got[0] - want[0]) <= REL * abs(want[0])
        assert tuple(st) == tuple(int(v) for v in wst)
    # the reference never terminates on these; the library reports them
    with pytest.raises(CsmError) as e:
        m.GetTransformation(empty, a, -0.1, 0.1)
    assert e.value.code == -4
    with pytest.raises(CsmError) as e:
        m.GetTransformation(a, a, 0.3, 0.3)
    assert e.value.code == -4


def test_rot_step_parameter_vs_oracle():
    """BASELINE config 4 uses a 0.25 degree step: an added parameter whose
    default reproduces the reference's 1 degree."""
    step = np.pi / 720
    pm = port.PortMatcher(30, 2, 0.3, 0.03, rot_step=step)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03, rot_step=step)
    a, b, _ = synth.make_pair(41, n_beams=360)
    got, st = m.GetTransformation(a, b, -0.3, 0.3, with_stats=True)
    want, wst = pm.match(a, b, -0.3, 0.3, with_stats=True)
    assert pose(got) == (want[1], want[2], want[3])
    assert tuple(st) == tuple(int(v) for v in wst)


def test_wide_window_vs_oracle():
    """+-10 m translation window (67x67 low-res translations), ragged cloud sizes."""
    pm = port.PortMatcher(30, 10, 0.3, 0.03)
    m = CorrelativeScanMatcher(30, 10, 0.3, 0.03)
    a, b, _ = synth.make_pair(42, trans_range=6.0, n_beams=200)
    got, st = m.GetTransformation(a[:150], b, -0.2, 0.2, with_stats=True)
    want, wst = pm.match(a[:150], b, -0.2, 0.2, with_stats=True)
    assert pose(got) == (want[1], want[2], want[3])
    assert tuple(st) == tuple(int(v) for v in wst)


def test_batch_equals_single_and_oracle():
    xy, offs = synth.make_scan_set(20, group=5, seed=3, n_beams=360)
    pa, pb = synth.make_pair_list(20, group=5)
    rng = np.random.default_rng(0)
    perm = rng.permutation(len(pa))  # callers may pass pairs in any order
    pa, pb = pa[perm], pb[perm]
    rmin = np.where(np.arange(len(pa)) % 2 == 0, -0.5, -0.3)
    rmax = np.where(np.arange(len(pa)) % 2 == 0, 0.5, 0.4)
    # max_tables=3 forces several table waves
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03, max_tables=3)
    m.upload_scans(xy, offs)
    score, x, y, th, st = m.match_batch(pa, pb, rmin, rmax, with_stats=True)
    assert m.last_timings()["waves"] == 7
    pm = port.PortMatcher(30, 2, 0.3, 0.03)
    for k in range(0, len(pa), 7):
        a = xy[offs[pa[k]]:offs[pa[k] + 1]]
        b = xy[offs[pb[k]]:offs[pb[k] + 1]]
        want, wst = pm.match(a, b, rmin[k], rmax[k], with_stats=True)
        assert (x[k], y[k], th[k]) == (want[1], want[2], want[3]), k
        assert abs(score[k] - want[0]) <= REL * abs(want[0])
        assert tuple(st[k]) == tuple(int(v) for v in wst)
        single = m.GetTransformation(a, b, rmin[k], rmax[k])
        assert pose(single) == (x[k], y[k], th[k]) and single[0] == score[k]


def test_full_size_properties():
    """Size-independent checks at the benchmark's scan size: self-match is the
    identity cell, results are deterministic, batch order does not matter."""
    xy, offs = synth.make_scan_set(40, group=5, seed=9)
    pa, pb = synth.make_pair_list(40, group=5)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m.upload_scans(xy, offs)
    r1 = m.match_batch(pa, pb, -0.5, 0.5)
    r2 = m.match_batch(pa[::-1].copy(), pb[::-1].copy(), -0.5, 0.5)
    for u, v in zip(r1, r2):
        assert np.array_equal(u, v[::-1])
    self_pairs = np.nonzero(pa == pb)[0]
    assert len(self_pairs) == 40
    # a scan against its own table: best translation within one low-res cell of 0
    assert np.all(np.abs(r1[1][self_pairs]) < 0.3) and np.all(np.abs(r1[2][self_pairs]) < 0.3)
    assert np.all(np.abs(r1[3][self_pairs]) < 0.036)  # within two 1-degree steps of 0
    assert np.all(r1[0] <= 0.0) and np.all(r1[0] >= np.log(1e-10))
    cs1 = m.build_tables(np.arange(40, dtype=np.int32))
    cs2 = m.build_tables(np.arange(40, dtype=np.int32)[::-1].copy())
    assert cs1 > 0 and abs(cs1 - cs2) <= 1e-9 * cs1


@pytest.mark.gpu
@pytest.mark.parametrize("cfg,w", [((40, 2, 0.3, 0.03), 2666), ((40, 2, 0.3, 0.0125), 6400)])
def test_grid_wider_than_checkpoint_budget_vs_oracle(cfg, w):
    """Grids wider than the 2048 cells the small checkpointed x/y passes cover:
    2666 cells per line takes the large-capacity instantiation (up to 6144 cells, the
    one the solver's 0.01 m tables use), 6400 cells the stored variants
    (csm_tables.cu: xpass_stored_kernel / ypass_stored_kernel).  Tables bit-identical
    with margin -1, the match identical to the oracle's."""
    pm = port.PortMatcher(*cfg)
    a, b, _ = synth.make_pair(23)
    ph = pm.highres(b)
    pl = pm.lowres(ph)
    m = CorrelativeScanMatcher(*cfg, margin_cells=-1)
    assert m.w_high == w
    hi, lo = m.GetLookupTables(b)
    assert np.array_equal(hi.values, ph.values) and np.array_equal(lo.values, pl.values)
    m.close()
    m = CorrelativeScanMatcher(*cfg)
    got, stats = m.GetTransformation(a, b, -0.5, 0.5, with_stats=True)
    want, wstats = pm.match(a, b, -0.5, 0.5, with_stats=True)
    assert (got[1][0][0], got[1][0][1], got[1][1]) == (want[1], want[2], want[3])
    assert abs(got[0] - want[0]) <= REL * abs(want[0])
    assert tuple(stats) == tuple(int(v) for v in wstats)
    m.close()


@pytest.mark.gpu
def test_concurrent_callers_on_one_handle():
    """The reference object is called from OpenMP threads (CorrelativeScanMatcher.cc:381-386,
    bag_avg_matcher.cc:104-141): concurrent single-pair calls on one handle must give the
    results of the same calls made one after the other (the handle serialises them)."""
    import threading
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    pairs = [synth.make_pair(40 + k, n_beams=240)[:2] for k in range(12)]
    serial = [m.GetTransformation(a, b, -0.5, 0.5) for a, b in pairs]
    out = [None] * len(pairs)
    errs = []

    def work(k):
        try:
            for _ in range(3):
                out[k] = m.GetTransformation(pairs[k][0], pairs[k][1], -0.5, 0.5)
                hi = m.GetLookupTableHighRes(pairs[k][1])
                assert hi.values.max() == 1.0
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=work, args=(k,)) for k in range(len(pairs))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    served, batches = m.single_queue_stats()
    assert served == len(pairs) * 4 and 1 <= batches <= served  # combined, never dropped
    for got, want in zip(out, serial):
        assert got[0] == want[0]
        assert got[1][0][0] == want[1][0][0] and got[1][0][1] == want[1][0][1] and got[1][1] == want[1][1]
    m.close()


@pytest.mark.gpu
def test_two_devices_in_one_process():
    """Handles on two devices of one process, used from two threads at once.  The
    >48 KB dynamic shared-memory opt-in of the low-res scorer, the refine kernels and
    the NN kernel is a per-device attribute (csm_common.cuh SmemOptIn): the second
    device's launches must not be refused, and both devices must return what one does."""
    import threading
    import torch
    from nautilus_b200 import pose_transform
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    xy, offs = synth.make_scan_set(20, group=5, seed=11, n_beams=360)
    pa, pb = synth.make_pair_list(20, group=5)
    tr = np.stack([pose_transform((0.1, 0.0, 0.02), (0.0, 0.05, 0.0))] * len(pa))
    out, errs = {}, []

    def work(dev):
        try:
            m = CorrelativeScanMatcher(30, 2, 0.3, 0.03, device=dev)
            m.upload_scans(xy, offs)
            r = m.match_batch(pa, pb, -0.5, 0.5, with_stats=True)
            m.nn_build(0.25)
            idx, _ = m.nn_match_batch(pa, pb, tr, 0.25)
            single = m.GetTransformation(xy[offs[0]:offs[1]], xy[offs[1]:offs[2]], -0.5, 0.5)
            out[dev] = (r, idx, single)
            m.close()
        except Exception as e:  # noqa: BLE001
            errs.append((dev, e))

    ts = [threading.Thread(target=work, args=(d,)) for d in (1, 0)]  # device 1 first
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    for u, v in zip(out[0][0], out[1][0]):
        assert np.array_equal(u, v)
    assert np.array_equal(out[0][1], out[1][1])
    assert out[0][2] == out[1][2]
