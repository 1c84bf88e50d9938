"""Golden fixtures for the shape of BASELINE configs[3] (wide search): translation
window +-10 m (67 x 67 low-res translations) and the unrestricted rotation +-pi.
The reference hard-codes a 1 degree step (CorrelativeScanMatcher.cc:92-93), so the
fixture pins 1 degree against the UNMODIFIED reference (oracle/_ref); the 0.25 degree
step of configs[3] is an added parameter and is checked against the oracle port in
tests/test_gpu_wide.py.  Run in the build container only:
    python tests/golden/make_wide_golden.py"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from nautilus_b200 import synth  # noqa: E402
from oracle import build, ref  # noqa: E402

PARAMS = (30.0, 10.0, 0.3, 0.03)


def wide_pairs():
    """Four small pairs: same room with a large offset; a cross-room pair (no overlap:
    flat landscape, ~140 pops); a scan against a 0.3 m wall stub (359 pops at 1 degree:
    more than the 255 windows the speculative refine lists per pair, so the replay
    carries on with the serial loop); a target out of reach (every candidate scores
    the floor: all 4489 windows tie and are popped in slot order)."""
    a0, b0, _ = synth.make_pair(70, trans_range=8.0, max_rot=2.5, n_beams=96)
    a1, _, _ = synth.make_pair(71, n_beams=96)
    _, b1, _ = synth.make_pair(72, n_beams=96)
    a2, _, _ = synth.make_pair(238, n_beams=48)
    rng = np.random.default_rng(38)
    c = rng.uniform(-3, 3, 2)
    b2 = (c + np.stack([np.linspace(0, 0.3, 6), np.zeros(6)], 1)
          + rng.normal(0, 0.01, (6, 2))).astype(np.float32)
    a3 = np.ascontiguousarray(a2[::3])
    b3 = np.asarray([[27.0, 27.5], [27.1, 27.5], [-28.0, 26.0]], np.float32)
    return [(a0, b0), (a1, b1), (a2, b2), (a3, b3)]


def main():
    assert build.build_ref() is not None, "/root/reference not available"
    rm = ref.RefMatcher(*PARAMS)
    out = {"params": np.asarray(PARAMS)}
    rows = []
    for k, (a, b) in enumerate(wide_pairs()):
        t0 = time.time()
        r = rm.match(a, b, -np.pi, np.pi)
        print(k, len(a), len(b), f"{time.time() - t0:.1f} s", r)
        out[f"a_{k}"], out[f"b_{k}"] = a, b
        rows.append([r[0], r[1], r[2], r[3]])
    out["match_1deg"] = np.asarray(rows, np.float64)
    path = os.path.join(HERE, "wide_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path))


if __name__ == "__main__":
    main()
