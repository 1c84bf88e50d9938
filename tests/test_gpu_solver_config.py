"""GPU parity for the ONE matcher configuration the Nautilus node itself constructs:
CorrelativeScanMatcher(30, 2, 0.3, 0.01) through the 5-argument overload with a
90 degree window (src/optimization/solver.cc:56, 633-638).  W = 6000 cells per
line, 30 x 30 hi-res candidates per popped window.

Fixtures: tests/golden/solver_golden.npz, generated from the unmodified reference
(oracle/_ref) by tests/golden/make_solver_golden.py; counters against the oracle
port, which the CPU suite pins to the same fixtures.

Bars: pose bit-exact, score <= 1e-5 relative, tables bit-identical with
margin_cells = -1 (sha256 of the fp64 bytes) and <= 1e-5 relative with the default
crop."""
import hashlib
import os

import numpy as np
import pytest

from conftest import ROOT
from nautilus_b200 import CorrelativeScanMatcher
from oracle import port

pytestmark = pytest.mark.gpu
REL = 1e-5


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(ROOT, "tests", "golden", "solver_golden.npz"))


@pytest.fixture(scope="module")
def matcher(g):
    m = CorrelativeScanMatcher(*g["params"])
    yield m
    m.close()


def test_solver_table_bit_identical_uncropped(g):
    m = CorrelativeScanMatcher(*g["params"], margin_cells=-1)
    assert (m.w_high, m.w_low) == (6000, 200)
    for k in (0, 2):
        hi, lo = m.GetLookupTables(g[f"b_{k}"])
        assert hashlib.sha256(hi.values.tobytes()).digest() == bytes(g[f"hi_sha_{k}"]), k
        assert np.array_equal(lo.values, g[f"lo_{k}"]), k
    m.close()


def test_solver_table_cropped_within_bar(g, matcher):
    for k in (0, 2):
        hi, lo = matcher.GetLookupTables(g[f"b_{k}"])
        got, want = hi.values[g[f"hi_ys_{k}"], g[f"hi_xs_{k}"]], g[f"hi_vals_{k}"]
        assert float((np.abs(got - want) / np.maximum(np.abs(want), 1e-10)).max()) <= REL
        assert float((np.abs(lo.values - g[f"lo_{k}"]) / np.maximum(g[f"lo_{k}"], 1e-10)).max()) <= REL
        assert abs(hi.values.sum() - float(g[f"hi_sum_{k}"])) <= 1e-9 * float(g[f"hi_sum_{k}"])
        assert hi.values.max() == 1.0


def test_solver_restricted_matches_vs_reference(g, matcher):
    """GetRelativeTransform's call (solver.cc:633-638) on five pairs, two of them with
    a window that wraps past pi."""
    restriction = float(g["restriction"])
    for k, (rot_a, rot_b, s, x, y, th) in enumerate(g["match"]):
        r = matcher.GetTransformationRestricted(g[f"a_{k}"], g[f"b_{k}"], rot_a, rot_b, restriction)
        assert (r[1][0][0], r[1][0][1], r[1][1]) == (np.float32(x), np.float32(y), np.float32(th)), (k, r)
        assert abs(r[0] - s) <= REL * abs(s), k


def test_solver_counters_vs_oracle(g, matcher):
    pm = port.PortMatcher(*g["params"])
    for k in (1, 3):
        rot_a, rot_b = g["match"][k][:2]
        lo, hi = matcher.rotation_window(rot_a, rot_b, float(g["restriction"]))
        assert (lo, hi) == pm.rotation_window(rot_a, rot_b, float(g["restriction"]))
        got, st = matcher.GetTransformation(g[f"a_{k}"], g[f"b_{k}"], lo, hi, with_stats=True)
        want, wst = pm.match(g[f"a_{k}"], g[f"b_{k}"], lo, hi, with_stats=True)
        assert (got[1][0][0], got[1][0][1], got[1][1]) == (want[1], want[2], want[3])
        assert tuple(st) == tuple(int(v) for v in wst), k


def test_solver_batch_equals_single(g, matcher):
    """The batched form a patched AddLCConstraints would call: all five pairs in one
    nautilus_csm_match_batch with per-pair windows."""
    clouds = [g[f"a_{k}"] for k in range(5)] + [g[f"b_{k}"] for k in range(5)]
    offs = np.zeros(11, np.int64)
    offs[1:] = np.cumsum([len(c) for c in clouds])
    matcher.upload_scans(np.concatenate(clouds), offs)
    wins = [matcher.rotation_window(r[0], r[1], float(g["restriction"])) for r in g["match"]]
    score, x, y, th = matcher.match_batch(np.arange(5), np.arange(5, 10),
                                          [w[0] for w in wins], [w[1] for w in wins])
    for k, row in enumerate(g["match"]):
        assert (x[k], y[k], th[k]) == (np.float32(row[3]), np.float32(row[4]), np.float32(row[5])), k
        assert abs(score[k] - row[2]) <= REL * abs(row[2])


def test_solver_regular_cell_rule(matcher):
    n, bad_low, bad_high = matcher.regular_check(300000, seed=11)
    assert n > 1000000 and bad_low == 0 and bad_high == 0
