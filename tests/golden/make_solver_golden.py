"""Golden fixtures for the ONE matcher configuration the Nautilus node itself
constructs: CorrelativeScanMatcher(30, 2, 0.3, 0.01) called through the 5-argument
overload with a 90 degree window (src/optimization/solver.cc:56, 633-638).

Generated from the UNMODIFIED reference (oracle/_ref, built from /root/reference by
oracle/build.py).  Run in the build container only:
    python tests/golden/make_solver_golden.py
W = 6000: a dense table is 288 MB, so the fixture holds the tables as digests
(sha256 of the raw fp64 bytes, which pins bit-identity) plus a sample of cells
(which pins the 1e-5 bar for cropped tables)."""
import hashlib
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from nautilus_b200 import synth  # noqa: E402
from oracle import build, ref  # noqa: E402

PARAMS = (30.0, 2.0, 0.3, 0.01)
RESTRICTION = np.pi / 2  # math_util::DegToRad(90), solver.cc:638


def rotate(pts, ang):
    c, s = np.cos(ang), np.sin(ang)
    return np.ascontiguousarray(
        np.stack([c * pts[:, 0] - s * pts[:, 1], s * pts[:, 0] + c * pts[:, 1]], 1), np.float32)


def sample_cells(values, seed, n=6000):
    """Cell indices worth pinning: every cell above 1e-3 on a stride, plus random ones."""
    rng = np.random.default_rng(seed)
    h, w = values.shape
    ys, xs = np.nonzero(values > 1e-8)
    pick = rng.choice(len(ys), size=min(n, len(ys)), replace=False)
    ry, rx = rng.integers(0, h, 500), rng.integers(0, w, 500)
    return (np.concatenate([ys[pick], ry]).astype(np.int32),
            np.concatenate([xs[pick], rx]).astype(np.int32))


def main():
    assert build.build_ref() is not None, "/root/reference not available"
    rm = ref.RefMatcher(*PARAMS)
    out = {"params": np.asarray(PARAMS), "restriction": np.asarray(RESTRICTION)}
    rows = []
    # (seed, beams, extra rotation of cloud a, offset of the window centre from the truth)
    cases = [(60, 1080, 0.0, 0.1), (61, 1080, 0.0, -0.3), (62, 360, 0.0, 0.6),
             (63, 360, 3.0, 0.05),   # rot_a - rot_b near pi: the window wraps (+2 pi branch, cc:255-257)
             (64, 720, -2.9, -0.2)]  # wraps on the other side
    for k, (seed, beams, extra, off) in enumerate(cases):
        a, b, truth = synth.make_pair(seed, n_beams=beams)
        a = rotate(a, -extra)  # the transform that moves a onto b now rotates by truth + extra
        rot_b = 0.37 * (k + 1)
        rot_a = rot_b + truth[2] + extra + off
        t0 = time.time()
        r = rm.match_restricted(a, b, rot_a, rot_b, RESTRICTION)
        print(f"pair {k}: {time.time() - t0:.1f} s", r, "truth", truth[0], truth[1], truth[2] + extra)
        out[f"a_{k}"], out[f"b_{k}"] = a, b
        rows.append([rot_a, rot_b, r[0], r[1], r[2], r[3]])
    out["match"] = np.asarray(rows, np.float64)
    # tables of two clouds: digest + sampled cells
    for k in (0, 2):
        hi = rm.highres(out[f"b_{k}"])
        lo = rm.lowres(hi)
        hv, lv = hi.values(), lo.values()
        assert hv.shape == (6000, 6000) and lv.shape == (200, 200)
        out[f"hi_sha_{k}"] = np.frombuffer(hashlib.sha256(hv.tobytes()).digest(), np.uint8)
        out[f"lo_{k}"] = lv
        ys, xs = sample_cells(hv, k)
        out[f"hi_ys_{k}"], out[f"hi_xs_{k}"], out[f"hi_vals_{k}"] = ys, xs, hv[ys, xs]
        out[f"hi_sum_{k}"] = np.asarray(hv.sum())
    path = os.path.join(HERE, "solver_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path))


if __name__ == "__main__":
    main()
