"""GPU parity on the full shape of BASELINE configs[3]: translation window +-10 m
(67 x 67 low-res translations), rotation +-pi, at the reference's 1 degree step against
fixtures from the UNMODIFIED reference (tests/golden/wide_golden.npz) and at the
0.25 degree step of configs[3] (an added parameter; the reference cannot step it,
CorrelativeScanMatcher.cc:92-93) against the oracle port.

The 67 x 67 grid is where the speculative refine (csm_search.cu, K6) leaves its fully
listed path: pairs 2 and 3 pop more than the 255 windows listed per pair, so the
replay hands over to the serial loop."""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

from conftest import ROOT
from nautilus_b200 import CorrelativeScanMatcher
from oracle import port

pytestmark = pytest.mark.gpu
REL = 1e-5


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(ROOT, "tests", "golden", "wide_golden.npz"))


def test_wide_1deg_vs_reference(g):
    m = CorrelativeScanMatcher(*g["params"])
    pops = []
    for k, (s, x, y, th) in enumerate(g["match_1deg"]):
        r, st = m.GetTransformation(g[f"a_{k}"], g[f"b_{k}"], -np.pi, np.pi, with_stats=True)
        assert (r[1][0][0], r[1][0][1], r[1][1]) == (np.float32(x), np.float32(y), np.float32(th)), (k, r)
        assert abs(r[0] - s) <= REL * abs(s), k
        assert st[0] == 361 * 67 * 67
        pops.append(st[2])
    assert pops[2] > 256 and pops[3] == 67 * 67  # beyond the speculative list: serial tail
    m.close()


def test_wide_quarter_degree_vs_oracle(g):
    """configs[3] as a whole: T = 10 m, rot_step = pi/720, +-pi; one batch call."""
    step = np.pi / 720
    pm = port.PortMatcher(*g["params"], rot_step=step)
    n = len(g["match_1deg"])
    clouds = [g[f"a_{k}"] for k in range(n)] + [g[f"b_{k}"] for k in range(n)]  # (npz reads are not thread-safe)
    with ThreadPoolExecutor(max_workers=n) as ex:  # ctypes releases the GIL
        want = list(ex.map(lambda k: pm.match(clouds[k], clouds[n + k], -np.pi, np.pi,
                                              with_stats=True), range(n)))
    offs = np.zeros(2 * n + 1, np.int64)
    offs[1:] = np.cumsum([len(c) for c in clouds])
    m = CorrelativeScanMatcher(*g["params"], rot_step=step)
    m.upload_scans(np.concatenate(clouds), offs)
    score, x, y, th, st = m.match_batch(np.arange(n), np.arange(n, 2 * n), -np.pi, np.pi,
                                        with_stats=True)
    for k, (w, wst) in enumerate(want):
        assert (x[k], y[k], th[k]) == (w[1], w[2], w[3]), (k, w)
        assert abs(score[k] - w[0]) <= REL * abs(w[0])
        assert tuple(st[k]) == tuple(int(v) for v in wst), k
    assert st[0, 0] == 1440 * 67 * 67 and st[2, 2] > 256
    m.close()
