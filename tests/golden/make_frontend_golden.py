"""Generates tests/golden/frontend_golden.npz from the UNMODIFIED reference (oracle/_ref:
third_party/csm/src/pointcloud_helpers.cc, src/input/pointcloud_helpers.cc and
src/loop_closure/lc_candidate_filter.cc compiled where they lie): outputs of both
LaserScanToPointCloud variants on seeded range scans, ComputeScatterMatrixScore of the
resulting clouds and GetLCCandidates over a seeded trajectory.
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
# the variant the Nautilus node itself calls (src/input/pointcloud_helpers.cc:28-48)
clouds = [ref.nautilus_laser_to_points(r[s], sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN,
                                       sr.MAX_RANGE) for s in range(len(r))]
offs = np.zeros(len(r) + 1, np.int64)
offs[1:] = np.cumsum([len(c) for c in clouds])
out["nautilus_xy"] = np.concatenate(clouds, axis=0)
out["nautilus_offsets"] = offs
# scatter score + keyframe selection (lc_candidate_filter.cc:34-81) over clouds of varied
# spread: real scans (elongated rooms score low) and blobs / ellipses that pass 0.70
rng = np.random.default_rng(9)
lc_clouds = []
for k in range(48):
    if k % 3 == 0:
        lc_clouds.append(clouds[k % len(clouds)] if len(clouds[k % len(clouds)]) > 2 else clouds[0])
    else:
        ax = rng.uniform(0.5, 1.0)
        lc_clouds.append((rng.normal(size=(int(rng.integers(50, 400)), 2)) * [3.0, 3.0 * ax])
                         .astype(np.float32))
lc_offs = np.zeros(len(lc_clouds) + 1, np.int64)
lc_offs[1:] = np.cumsum([len(c) for c in lc_clouds])
lc_xy = np.concatenate(lc_clouds, axis=0)
lc_poses = np.concatenate([np.cumsum(rng.normal(0, 1.6, size=(48, 2)), axis=0),
                           rng.uniform(-3, 3, size=(48, 1))], axis=1)
out["lc_xy"], out["lc_offsets"], out["lc_poses"] = lc_xy, lc_offs, lc_poses
out["lc_scores"] = np.asarray([ref.scatter_score(c) for c in lc_clouds])
out["lc_candidates"] = ref.lc_candidates(lc_xy, lc_offs, lc_poses)
assert 3 <= len(out["lc_candidates"]) < 40, out["lc_candidates"]
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "frontend_golden.npz"), **out)
print({k: v.shape for k, v in out.items()})
