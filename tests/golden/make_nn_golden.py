"""Generates tests/golden/nn_golden.npz from the UNMODIFIED reference KD-tree
(oracle/_ref, /root/reference/src/util/kdtree.cc compiled by oracle/build.py).
Run in the build container only:
    python tests/golden/make_nn_golden.py
The reference has no test or fixture for its nearest-neighbour path; these
vectors are outputs of the reference itself on seeded inputs."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from nautilus_b200 import synth  # noqa: E402
from oracle import build, ref  # noqa: E402


def perturbed_poses(poses, rng, sigma_xy=0.08, sigma_th=0.02):
    """Odometry-like pose estimates around the true scan poses."""
    p = np.array(poses, dtype=np.float64)
    p[:, :2] += rng.normal(0, sigma_xy, size=(len(p), 2))
    p[:, 2] += rng.normal(0, sigma_th, size=len(p))
    return p


def adversarial_clouds():
    """Clouds that force exact float ties, duplicate points, axis-aligned
    separations and the single-child corner of the reference's tree walk."""
    rng = np.random.default_rng(42)
    g = np.stack(np.meshgrid(np.arange(-8, 9), np.arange(-6, 7)), -1).reshape(-1, 2) * 0.125
    lattice = np.ascontiguousarray(g, np.float32)                       # exact binary fractions
    dup = np.repeat(rng.uniform(-2, 2, size=(60, 2)).astype(np.float32), 3, axis=0)
    rng.shuffle(dup)
    line = np.stack([np.linspace(-3, 3, 97), np.zeros(97)], 1).astype(np.float32)  # y identical
    pairs2 = rng.uniform(-3, 3, size=(2, 2)).astype(np.float32)         # a two-point cloud
    one = np.asarray([[0.5, -0.25]], np.float32)
    sparse = rng.uniform(-5, 5, size=(300, 2)).astype(np.float32)
    return [lattice, dup, line, pairs2, one, sparse, np.zeros((0, 2), np.float32)]


def main():
    assert build.build_ref() is not None, "/root/reference not available"
    out = {}
    # 1. synthetic scans, window-style pairs, default outlier threshold
    xy, offs, poses = synth.make_scan_set(20, group=5, seed=11, n_beams=360, return_poses=True)
    rng = np.random.default_rng(5)
    est = perturbed_poses(poses, rng)
    src, tgt = synth.make_pair_list(20, group=5)
    out["scan_xy"], out["scan_offsets"], out["scan_poses"] = xy, offs, est
    out["scan_src"], out["scan_tgt"] = src, tgt
    # 2. adversarial clouds, every ordered pair, lattice-aligned and random transforms
    clouds = adversarial_clouds()
    axy = np.concatenate(clouds, 0)
    aoffs = np.zeros(len(clouds) + 1, np.int64)
    aoffs[1:] = np.cumsum([len(c) for c in clouds])
    asrc, atgt, am = [], [], []
    for s in range(len(clouds)):
        for t in range(len(clouds)):
            for m in ([1, 0, 0, 0, 1, 0], [1, 0, 0.0625, 0, 1, -0.0625], [0, -1, 0.125, 1, 0, 0],
                      [0.96, -0.28, 0.03, 0.28, 0.96, -0.02]):
                asrc.append(s); atgt.append(t); am.append(m)
    out["adv_xy"], out["adv_offsets"] = axy, aoffs
    out["adv_src"], out["adv_tgt"] = np.asarray(asrc, np.int32), np.asarray(atgt, np.int32)
    out["adv_m"] = np.asarray(am, np.float32)
    out["thresholds"] = np.asarray([0.25, 0.05, 0.7, 1e-5])  # 0.7f < 0.7, 0.05f > 0.05
    return out, (xy, offs, est, src, tgt), (axy, aoffs, asrc, atgt, am)


def run():
    out, scan, adv = main()
    from nautilus_b200.point_matching import pose_transform
    xy, offs, est, src, tgt = scan
    m = np.stack([pose_transform(est[s], est[t]) for s, t in zip(src, tgt)])
    out["scan_m"] = m
    for k, thr in enumerate(out["thresholds"]):
        _, idx, o, dist = ref.nn_match_batch(xy, offs, src, tgt, m, thr, with_dist=True)
        out[f"scan_idx_{k}"], out[f"scan_dist_{k}"], out["scan_out_offsets"] = idx, dist, o
        _, idx, o, dist = ref.nn_match_batch(out["adv_xy"], out["adv_offsets"], out["adv_src"],
                                             out["adv_tgt"], out["adv_m"], thr, with_dist=True)
        out[f"adv_idx_{k}"], out[f"adv_dist_{k}"], out["adv_out_offsets"] = idx, dist, o
    # GetPointNormal's lookup: every point in its own tree with threshold 1e-5
    for name in ("scan", "adv"):
        pts, o = out[f"{name}_xy"], out[f"{name}_offsets"]
        self_idx = np.zeros(len(pts), np.int32)
        for s in range(len(o) - 1):
            c = pts[o[s]:o[s + 1]]
            if len(c):
                d, i = ref.RefKDTree(c).nearest(c, 1e-5)
                assert (d < 1e-5).all()
                self_idx[o[s]:o[s + 1]] = i
        out[f"{name}_self_idx"] = self_idx
    np.savez_compressed(os.path.join(HERE, "nn_golden.npz"), **out)
    print("wrote nn_golden.npz:", {k: v.shape for k, v in out.items() if k.endswith("idx_0")})


if __name__ == "__main__":
    run()
