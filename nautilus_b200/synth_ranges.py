"""Seeded synthetic LaserScan range batches (the shape of a bag's scans before
LaserScanToPointCloud): smooth wall-like profiles with out-of-range, too-near, NaN
and inf returns.  Shared by tests and bench side-configs."""
from __future__ import annotations

import numpy as np

N_BEAMS = 1080
ANGLE_MIN = np.float32(-3.0 * np.pi / 4.0)
ANGLE_INC = np.float32(1.5 * np.pi / N_BEAMS)
ANGLE_MAX = np.float32(ANGLE_MIN + ANGLE_INC * np.float32(N_BEAMS))
RANGE_MIN = np.float32(0.05)
MAX_RANGE = 30.0


def make_ranges(n_scans: int, n_beams: int = N_BEAMS, seed: int = 0, dirty: bool = True) -> np.ndarray:
    rng = np.random.default_rng(4242 + seed)
    base = rng.uniform(0.5, 20.0, size=(n_scans, 1))
    r = base + np.cumsum(rng.normal(0.0, 0.03, size=(n_scans, n_beams)), axis=1)
    # a few depth jumps per scan (doorways): these are what the glancing filter drops
    for s in range(n_scans):
        for at in rng.integers(0, max(n_beams, 1), size=4):
            r[s, at:] += rng.uniform(-3.0, 3.0)
    r = np.abs(r).astype(np.float32)
    if dirty:
        u = rng.random((n_scans, n_beams))
        r[u < 0.05] = np.float32(40.0)
        r[(u >= 0.05) & (u < 0.07)] = np.float32(0.01)
        r[(u >= 0.07) & (u < 0.08)] = np.nan
        r[(u >= 0.08) & (u < 0.09)] = np.inf
    return np.ascontiguousarray(r)
