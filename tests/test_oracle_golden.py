"""CPU tests: the oracle port (oracle/csm_oracle.c) against
  (1) the reference's own 11 unit tests
      (third_party/csm/test/correlative_scan_matcher_tests.cc:12-106),
  (2) fixtures generated from the unmodified reference (tests/golden),
  (3) the unmodified reference itself when oracle/_ref is present.
Bit-exact throughout: integer/index work and fp64 tables with identical
operation order."""
import numpy as np
import pytest

from nautilus_b200 import synth
from oracle import port


def T(range_, res):
    return port.PortTable(range_, res)


# ---- (1) the reference's unit tests, restated one to one ---------------------
def test_set_single_point():
    t = T(4, 1)
    t.set(0, 0, 0.5)
    assert t.get(0, 0) == 0.5
    assert t.get(.5, .5) == 0.5
    assert t.get(-.5, -.5) < 0.5


def test_overwrite_point():
    t = T(4, 1)
    t.set(3.9, 3.9, 0.01)
    assert t.get(3.9, 3.9) == 0.01
    t.set(-3.9, -3.9, 0.02)
    assert t.get(-3.9, -3.9) == 0.02


def test_nearby_points():
    t = T(4, 1)
    t.set(-3.9, -3.9, 0.01)
    t.set(-3.0, -3.0, 0.275)
    assert t.get(-3.51, -3.51) == 0.01
    assert t.get(-3.49, -3.49) == 0.01
    assert t.get(-2.99, -2.99) == 0.275


def test_boundary_edge():
    t = T(4, 1)
    t.set(-4, -4, 0.23)
    assert t.get(-4, -4) == 0.23
    assert t.get(-4, -3.99) == 0.23
    t.set(4, -4, 0.24)
    assert t.get(3.99, -4) < 0.24
    t.set(3.99, -4, 0.24)
    assert t.get(3.99, -3.99) == 0.24


def test_max_area():
    t = T(4, 1)
    for (x, y, v) in [(-3.5, -3, 0.2), (-3, 3, 0.1), (2, -4, 0.4), (2, 2, 0.5), (4, 0, 0.6)]:
        t.set(x, y, v)
    assert t.max_area(-4, -3, 3, 1) == 0.2


def test_out_of_bounds():
    t = T(4, 1)
    t.set(-4, -5, 0.5)
    assert t.max_area(-4, -4, 4, 4) < 0.5


def test_create_low_res_simple():
    t = T(4, 1)
    t.set(0, 0, 1)
    low = port.PortMatcher(4, 4, 1, 1).lowres(t)
    assert low.get(0, 0) == 1


def test_create_low_res():
    t = T(4, 0.5)
    t.set(0, 0.1, 0.5)
    t.set(0, 0.7, 0.56)
    low = port.PortMatcher(4, 4, 1, 0.5).lowres(t)
    assert low.get(0, 0) == 0.56


def test_low_res_inexpressible_resolution():
    t = T(0.3, 0.03)
    t.set(0, 0, 0.5)
    t.set(0, 0.031, 0.56)
    low = port.PortMatcher(0.3, 0.3, 0.3, 0.03).lowres(t)
    assert (low.width, low.height) == (2, 2)
    assert low.get(0, 0) == 0.56


def test_low_res_inexpressible_resolution_second():
    t = T(0.3, 0.03)
    t.set(0, 0, 0.5)
    t.set(0, -0.031, 0.56)
    t.set(-0.061, -0.031, 0.56)
    t.set(-0.031, -0.031, 0.75)
    low = port.PortMatcher(0.3, 0.3, 0.3, 0.03).lowres(t)
    assert (low.width, low.height) == (2, 2)
    assert low.get(0, 0) == 0.5
    assert low.get(-0.03, -0.03) == 0.75
    assert low.get(-0.03, 0.03) == 1e-10
    assert low.get(0.0, -0.03) == 0.56


# ---- (2) fixtures from the unmodified reference ------------------------------
def test_tables_match_golden(golden):
    pm = port.PortMatcher(*golden["small_params"])
    for k in range(4):
        hi = pm.highres(golden[f"small_cloud_{k}"])
        assert np.array_equal(hi.values, golden[f"small_hi_{k}"]), k
        lo = pm.lowres(hi)
        assert np.array_equal(lo.values, golden[f"small_lo_{k}"]), k


def test_matches_match_golden(golden):
    pm = port.PortMatcher(*golden["small_params"])
    for k, row in enumerate(golden["small_match"]):
        a, b = golden[f"small_a_{k}"], golden[f"small_b_{k}"]
        r1 = pm.match(a, b, -0.5, 0.5)
        r2 = pm.match(a, b)
        got = [r1[0], r1[1], r1[2], r1[3], r2[0], r2[1], r2[2], r2[3]]
        assert np.array_equal(np.asarray(got, np.float64), row), (k, got, row)


def test_std_matches_match_golden(golden):
    pm = port.PortMatcher(*golden["std_params"])
    for k, row in enumerate(golden["std_match"]):
        r = pm.match(golden[f"std_a_{k}"], golden[f"std_b_{k}"], -0.5, 0.5)
        assert np.array_equal(np.asarray(r, np.float64), row), (k, r, row)


def test_restricted_window_golden(golden):
    pm = port.PortMatcher(*golden["small_params"])
    a, b = golden["small_a_0"], golden["small_b_0"]
    for ra, rb, rr, s, x, y, th in golden["small_restricted"]:
        lo, hi = pm.rotation_window(ra, rb, rr)
        r = pm.match(a, b, lo, hi)
        assert np.array_equal(np.asarray(r, np.float64), [s, x, y, th])


# ---- (3) live against the unmodified reference -------------------------------
def test_port_equals_reference_live(ref_lib):
    rm = ref_lib.RefMatcher(30, 2, 0.3, 0.03)
    pm = port.PortMatcher(30, 2, 0.3, 0.03)
    a, b, _ = synth.make_pair(11, n_beams=270)
    rh = rm.highres(b)
    ph = pm.highres(b)
    assert np.array_equal(rh.values(), ph.values)
    assert np.array_equal(rm.lowres(rh).values(), pm.lowres(ph).values)
    assert rm.match(a, b, -0.5, 0.5) == pm.match(a, b, -0.5, 0.5)


def test_port_edge_inputs(ref_lib):
    """Empty target, points outside the grid, duplicated clouds (ties)."""
    rm = ref_lib.RefMatcher(4, 1, 0.5, 0.05)
    pm = port.PortMatcher(4, 1, 0.5, 0.05)
    a = np.asarray([[0.5, 0.5], [1.0, -1.0], [-2.0, 0.3]], np.float32)
    empty = np.zeros((0, 2), np.float32)
    assert rm.match(a, empty, -0.1, 0.1) == pm.match(a, empty, -0.1, 0.1)
    far = np.asarray([[10.0, 0.0], [0.0, -9.0], [0.2, 0.1]], np.float32)
    assert rm.match(a, far, -0.1, 0.1) == pm.match(a, far, -0.1, 0.1)
    assert rm.match(a, a, -0.2, 0.2) == pm.match(a, a, -0.2, 0.2)
    with pytest.raises(ValueError):
        pm.match(empty, a, -0.1, 0.1)


# ---- nearest-neighbour path: the restatement against reference outputs ----------
def test_nn_port_vs_reference_golden():
    """oracle/nn_oracle.cc reproduces the compiled reference KD-tree's matches
    (tests/golden/nn_golden.npz, generated by tests/golden/make_nn_golden.py)."""
    import os
    from oracle import nn_port
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "nn_golden.npz"))
    quirk = 0
    for name in ("scan", "adv"):
        for k, thr in enumerate(g["thresholds"]):
            idx, off, dist = nn_port.match_batch(g[f"{name}_xy"], g[f"{name}_offsets"],
                                                 g[f"{name}_src"], g[f"{name}_tgt"],
                                                 g[f"{name}_m"], float(thr))
            want = g[f"{name}_idx_{k}"]
            assert np.array_equal(off, g[f"{name}_out_offsets"])
            assert np.array_equal(idx, want), (name, thr)
            assert np.array_equal(dist[want >= 0], g[f"{name}_dist_{k}"][want >= 0])
        exact = nn_port.brute_force(g["scan_xy"], g["scan_offsets"], g["scan_src"][:20],
                                    g["scan_tgt"][:20], g["scan_m"][:20], 0.25)
        n = len(exact)
        quirk += int((exact != g["scan_idx_0"][:n]).sum())
    # the reference walk is not an exact nearest-neighbour search (single-child leaves)
    assert quirk > 0


def test_nn_port_vs_compiled_reference(ref_lib):
    from nautilus_b200 import synth
    from oracle import nn_port
    xy, offs, poses = synth.make_scan_set(10, group=5, seed=123, n_beams=360, return_poses=True)
    src, tgt = synth.make_pair_list(10, group=5)
    rng = np.random.default_rng(0)
    tr = np.tile(np.asarray([1, 0, 0, 0, 1, 0], np.float32), (len(src), 1))
    tr[:, [2, 5]] = rng.normal(0, 0.1, (len(src), 2))
    for thr in (0.25, 0.7):
        idx, off, _ = nn_port.match_batch(xy, offs, src, tgt, tr, thr)
        _, want, wo = ref_lib.nn_match_batch(xy, offs, src, tgt, tr, thr)
        assert np.array_equal(off, wo) and np.array_equal(idx, want)


# ---- the solver's own matcher: CSM(30, 2, 0.3, 0.01), 5-argument overload with a
# 90 degree window (src/optimization/solver.cc:56, 633-638); fixtures from oracle/_ref
@pytest.fixture(scope="module")
def solver_golden():
    import os
    from conftest import ROOT
    return np.load(os.path.join(ROOT, "tests", "golden", "solver_golden.npz"))


def test_port_solver_config_table_vs_golden(solver_golden):
    import hashlib
    g = solver_golden
    pm = port.PortMatcher(*g["params"])
    hi = pm.highres(g["b_2"])
    assert hi.values.shape == (6000, 6000)
    assert hashlib.sha256(hi.values.tobytes()).digest() == bytes(g["hi_sha_2"])
    assert np.array_equal(hi.values[g["hi_ys_2"], g["hi_xs_2"]], g["hi_vals_2"])
    assert np.array_equal(pm.lowres(hi).values, g["lo_2"])


@pytest.mark.parametrize("k", [2, 3])  # 3: the window wraps (rot_max += 2 pi, cc:255-257)
def test_port_solver_config_match_vs_golden(solver_golden, k):
    g = solver_golden
    pm = port.PortMatcher(*g["params"])
    rot_a, rot_b, s, x, y, th = g["match"][k]
    lo, hi = pm.rotation_window(rot_a, rot_b, float(g["restriction"]))
    r, st = pm.match(g[f"a_{k}"], g[f"b_{k}"], lo, hi, with_stats=True)
    assert r == (s, np.float32(x), np.float32(y), np.float32(th))
    assert st[1] == 900 * st[2]  # 30 x 30 hi-res candidates per popped window
