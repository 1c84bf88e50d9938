"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built
from /root/reference by oracle/build.py).  Run in the build container only:
    python tests/golden/make_golden.py
The fixtures pin the oracle port and the CUDA path on machines where
/root/reference does not exist (the GPU box)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from nautilus_b200 import synth  # noqa: E402
from oracle import build, ref  # noqa: E402


def small_cloud(seed, n=160, half=2.5):
    """Points on the walls of a small box + noise, inside a 4 m range table."""
    rng = np.random.default_rng(seed)
    t = rng.uniform(-half, half, size=n)
    side = rng.integers(0, 4, size=n)
    x = np.where(side == 0, -half, np.where(side == 1, half, t))
    y = np.where(side == 2, -half * 0.8, np.where(side == 3, half * 0.8, t * 0.8))
    y = np.where(side < 2, t * 0.8, y)
    pts = np.stack([x, y], 1) + rng.normal(0, 0.01, size=(n, 2))
    return np.ascontiguousarray(pts, np.float32)


def transform(pts, dx, dy, dth):
    c, s = np.cos(dth), np.sin(dth)
    out = np.stack([c * pts[:, 0] - s * pts[:, 1] + dx, s * pts[:, 0] + c * pts[:, 1] + dy], 1)
    return np.ascontiguousarray(out, np.float32)


def main():
    assert build.build_ref() is not None, "/root/reference not available"
    out = {}
    # 1. tables of a small matcher: CSM(4, 1, 0.5, 0.05) -> 160^2 hi, 16^2 low
    small = (4.0, 1.0, 0.5, 0.05)
    rm = ref.RefMatcher(*small)
    clouds = [small_cloud(s) for s in range(3)]
    # one cloud touching the grid edge (exercises Neumann seed / Triggs boundary)
    edge = small_cloud(7)
    edge[:20, 0] = 3.99
    edge[20:40, 1] = -4.0
    clouds.append(edge)
    for k, c in enumerate(clouds):
        hi = rm.highres(c)
        lo = rm.lowres(hi)
        out[f"small_cloud_{k}"] = c
        out[f"small_hi_{k}"] = hi.values()
        out[f"small_lo_{k}"] = lo.values()
    out["small_params"] = np.asarray(small)
    # 2. matches with the small matcher (full rotation and a window)
    res = []
    for k in range(6):
        b = small_cloud(100 + k, n=120)
        rng = np.random.default_rng(500 + k)
        a = transform(small_cloud(100 + k, n=120), *rng.uniform(-0.6, 0.6, 2), rng.uniform(-0.3, 0.3))
        out[f"small_a_{k}"], out[f"small_b_{k}"] = a, b
        r1 = rm.match(a, b, -0.5, 0.5)
        r2 = rm.match(a, b)
        res.append([r1[0], r1[1], r1[2], r1[3], r2[0], r2[1], r2[2], r2[3]])
    out["small_match"] = np.asarray(res, dtype=np.float64)  # float32 values are exact in float64
    # 3. default-window matches on synthetic 1080-beam scans (config 1 parameters)
    std = (30.0, 2.0, 0.3, 0.03)
    rs = ref.RefMatcher(*std)
    res = []
    for k in range(4):
        a, b, _ = synth.make_pair(k, n_beams=360)
        out[f"std_a_{k}"], out[f"std_b_{k}"] = a, b
        r = rs.match(a, b, -0.5, 0.5)
        res.append([r[0], r[1], r[2], r[3]])
    out["std_match"] = np.asarray(res, dtype=np.float64)
    out["std_params"] = np.asarray(std)
    # 4. 5-argument overload windows
    wins = []
    for ra, rb, rr in [(0.1, 0.3, np.pi / 2), (3.0, -3.0, 1.0), (-2.0, 2.5, 0.5), (0.0, 0.0, np.pi / 2)]:
        a, b = out["small_a_0"], out["small_b_0"]
        r = rm.match_restricted(a, b, ra, rb, rr)
        wins.append([ra, rb, rr, r[0], r[1], r[2], r[3]])
    out["small_restricted"] = np.asarray(wins, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "csm_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "csm_golden.npz"), os.path.getsize(os.path.join(HERE, "csm_golden.npz")))


if __name__ == "__main__":
    main()
