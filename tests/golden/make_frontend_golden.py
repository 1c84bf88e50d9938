"""Generates tests/golden/frontend_golden.npz from the UNMODIFIED reference
(oracle/_ref, third_party/csm/src/pointcloud_helpers.cc compiled where it lies):
outputs of pointcloud_helpers::LaserScanToPointCloud on seeded range scans.
Run in the build container: python tests/golden/make_frontend_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from nautilus_b200 import synth_ranges as sr  # noqa: E402
from oracle import build, ref  # noqa: E402

assert build.build_ref() is not None
r = sr.make_ranges(12, seed=5)
r[3, :] = np.float32(50.0)      # a scan with no valid return
r[4, 1:] = np.float32(50.0)     # a single point
out = {"ranges": r}
for tr in (0, 1):
    clouds = [ref.csm_laser_to_points(r[s], sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN,
                                      sr.MAX_RANGE, bool(tr)) for s in range(len(r))]
    offs = np.zeros(len(r) + 1, np.int64)
    offs[1:] = np.cumsum([len(c) for c in clouds])
    out[f"csm_t{tr}_xy"] = np.concatenate(clouds, axis=0)
    out[f"csm_t{tr}_offsets"] = offs
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "frontend_golden.npz"), **out)
print({k: v.shape for k, v in out.items()})
