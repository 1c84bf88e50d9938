"""Parity soak: pairs of several seeded batches matched on the B200 and by the
compiled reference (oracle/_ref) on the host cores; poses must be bit-identical,
scores within 1e-5 relative.  Writes one JSON line.  (Checker use of oracle/: this is
test tooling, like tests/ and bench.py's cpu_baseline leg.)

  python scripts/parity_soak.py [--seeds 4] [--pairs 1200]
"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from nautilus_b200 import CorrelativeScanMatcher


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=4)
    ap.add_argument("--pairs", type=int, default=1200, help="pairs checked per seed")
    ap.add_argument("--scans", type=int, default=400)
    ap.add_argument("--cross", action="store_true",
                    help="pair scans of DIFFERENT rooms (no overlap: flat score landscapes, many pops)")
    args = ap.parse_args()
    cores = bench.host_cores()
    m = CorrelativeScanMatcher(*bench.PARAMS, device=0)
    tot = bad_pose = 0
    worst = 0.0
    kind = None
    t0 = time.perf_counter()
    for seed in range(100, 100 + args.seeds):
        xy, offs, pa, pb = bench.make_batch(args.scans, seed=seed)
        if args.cross:   # same queries, targets drawn from anywhere in the set
            rng = np.random.default_rng(seed)
            pb = rng.integers(0, args.scans, size=len(pa)).astype(pb.dtype)
        m.upload_scans(xy, offs)
        score, x, y, th, st = m.match_batch(pa, pb, *bench.ROT_WINDOW, with_stats=True)
        idx = np.linspace(0, len(pa) - 1, min(args.pairs, len(pa))).astype(np.int64)
        secs, kind, (rs, rx, ry, rt) = bench.cpu_reference_run(xy, offs, pa, pb, idx, cores)
        same = (rx == x[idx]) & (ry == y[idx]) & (rt == th[idx])
        bad_pose += int((~same).sum())
        worst = max(worst, float(np.max(np.abs(score[idx] - rs) / np.abs(rs))))
        tot += len(idx)
    m.close()
    print(json.dumps({"check": "GPU match_batch vs reference CPU GetTransformation", "kind": kind,
                      "pairs": tot, "seeds": args.seeds, "cross_room_pairs": bool(args.cross), "pose_mismatches": bad_pose,
                      "worst_score_rel_err": worst, "cores": cores,
                      "seconds": round(time.perf_counter() - t0, 1)}), flush=True)
    sys.exit(1 if bad_pose or worst > 1e-5 else 0)


if __name__ == "__main__":
    main()
