"""A 500-pair slice of scripts/parity_soak.py under the driver's `-m gpu` run: pairs
of a seeded batch (same-room pairs: peaked landscapes; cross-room pairs: flat
landscapes with long pop chains through the speculative refine's replay) matched on
the B200 and by the UNMODIFIED reference (oracle/_ref, prebuilt, one pair per host
thread).  Poses bit-identical, scores <= 1e-5 relative."""
import os

import numpy as np
import pytest

from nautilus_b200 import CorrelativeScanMatcher, synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cross,n_check", [(False, 300), (True, 200)])
def test_soak_slice_vs_reference(ref_lib, cross, n_check):
    n_scans = 160
    xy, offs = synth.make_scan_set(n_scans, group=5, seed=321, n_beams=540)
    pa, pb = synth.make_pair_list(n_scans, group=5)
    if cross:  # same queries, targets drawn from anywhere in the set
        pb = np.random.default_rng(5).integers(0, n_scans, size=len(pa)).astype(pb.dtype)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m.upload_scans(xy, offs)
    score, x, y, th = m.match_batch(pa, pb, -0.5, 0.5)
    idx = np.linspace(0, len(pa) - 1, n_check).astype(np.int64)
    threads = len(os.sched_getaffinity(0))
    rm = ref_lib.RefMatcher(30, 2, 0.3, 0.03)
    _, rs, rx, ry, rt = rm.match_batch(xy, offs, pa[idx], pb[idx], np.full(n_check, -0.5),
                                       np.full(n_check, 0.5), threads=threads)
    bad = np.nonzero((rx != x[idx]) | (ry != y[idx]) | (rt != th[idx]))[0]
    assert len(bad) == 0, (idx[bad][:5], x[idx][bad][:5], rx[bad][:5])
    assert float(np.max(np.abs(score[idx] - rs) / np.abs(rs))) <= 1e-5
    m.close()
