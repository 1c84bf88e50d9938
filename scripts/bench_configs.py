"""Side measurements for the BASELINE configs that are not bench.py's headline
(SURVEY.md 8d): table build (configs[1]), wide search (configs[3]) and the
point-to-point NN path (configs[4]).  One JSON line per config, CUDA-event times
from the library, the reference's CPU code (oracle/_ref) timed beside it on a
bounded sample.  Usage (on a B200):

  python scripts/bench_configs.py [tables] [wide] [solver] [uncertainty] [nn] [frontend] [residuals] [--cpu]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from nautilus_b200 import CorrelativeScanMatcher, pose_transform, synth  # noqa: E402

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] \
    if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def cores():
    return len(os.sched_getaffinity(0))


def bench_tables(cpu):
    n = 4096
    xy, offs = synth.make_scan_set(n, group=8, seed=2)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m.upload_scans(xy, offs)
    ids = np.arange(n, dtype=np.int32)
    m.build_tables(ids[:256])
    ms = []
    for _ in range(3):
        cs = m.build_tables(ids)
        ms.append(m.last_timings()["ms_tables"])
    t = float(np.median(ms)) * 1e-3
    alg = 8.0 * (m.w_high ** 2 + m.w_low ** 2) * n
    line = {"config": "batch lookup-table rasterisation + blur, 4096 scans (BASELINE configs[1])",
            "metric": "tables_per_sec", "value": n / t, "ms": t * 1e3,
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_table": alg / n,
                         "achieved": alg / t / 1e9, "peak": PEAK, "unit": "GB/s",
                         "frac": alg / t / 1e9 / PEAK},
            "low_res_checksum": cs}
    if cpu:
        from oracle import ref
        k = 4 * cores()
        sub_off = offs[:k + 1] - offs[0]
        secs, rcs = ref.RefMatcher(30, 2, 0.3, 0.03).build_tables_batch(xy[:offs[k]], sub_off,
                                                                       threads=cores())
        gcs = m.build_tables(ids[:k])
        line["cpu_baseline"] = {"value": k / secs, "unit": "tables/s", "cores": cores(),
                                "kind": "reference", "sample": f"{k} tables, {secs:.1f} s",
                                "checksum_rel_diff": abs(gcs - rcs) / abs(rcs)}
    m.close()
    return line


def low_fp64_frac(st, offs, pa, ms_low, sm_mhz=1965.0):
    """K5's fraction of the fp64 add pipe, as bench.py states it: one add per (low-res
    candidate, point) over 148 SMs x 64 adds per clock."""
    pts = (offs[1:] - offs[:-1])[pa].astype(np.float64)
    adds = float((st[:, 0].astype(np.float64) * pts).sum())
    return adds / (ms_low * 1e-3) / (148 * 64 * sm_mhz * 1e6)


def bench_wide(cpu):
    n_scans, group = 40, 5
    xy, offs = synth.make_scan_set(n_scans, group=group, seed=4, trans_range=10.0, max_rot=1.5)
    pa, pb = synth.make_pair_list(n_scans, group=group)        # 200 pairs
    reps = 5                                                   # 1000 pair matches
    pa, pb = np.tile(pa, reps), np.tile(pb, reps)
    m = CorrelativeScanMatcher(30, 10, 0.3, 0.03, rot_step=np.pi / 720)
    m.upload_scans(xy, offs)
    m.match_batch(pa[:40], pb[:40], -np.pi, np.pi)
    t0 = time.perf_counter()
    s, x, y, th, st = m.match_batch(pa, pb, -np.pi, np.pi, with_stats=True)
    wall = time.perf_counter() - t0
    tm = m.last_timings()
    cand = int(st[:, 0].sum() + st[:, 1].sum())
    line = {"config": "wide search: +-pi at 0.25 deg, +-10 m at 0.03 m, 1000 pairs (BASELINE configs[3])",
            "metric": "candidate_transforms_scored_per_sec", "value": cand / (tm["ms_total"] * 1e-3),
            "pairs_per_sec": len(pa) / (tm["ms_total"] * 1e-3), "ms": tm["ms_total"],
            "ms_wall": wall * 1e3, "kernels_ms": {k: tm[k] for k in ("ms_tables", "ms_low", "ms_refine")},
            "low_candidates_per_pair": int(st[0, 0]), "mean_pops": float(st[:, 2].mean()),
            "low_score_kernel_fp64_add_frac": low_fp64_frac(st, offs, pa, tm["ms_low"])}
    if cpu:
        from oracle import ref
        k = min(cores(), 8)
        rm = ref.RefMatcher(30, 10, 0.3, 0.03)
        # the reference hard-codes 1 degree (cc:92): its wide search is timed at 1 degree
        # and compared through candidates/s, which is step-independent
        idx = np.arange(k)
        secs, rs, rx, ry, rt = rm.match_batch(xy, offs, pa[idx], pb[idx], np.full(k, -np.pi),
                                              np.full(k, np.pi), threads=k)
        m1 = CorrelativeScanMatcher(30, 10, 0.3, 0.03)
        m1.upload_scans(xy, offs)
        g = m1.match_batch(pa[idx], pb[idx], -np.pi, np.pi, with_stats=True)
        c1 = int(g[4][:, 0].sum() + g[4][:, 1].sum())
        line["cpu_baseline"] = {"value": c1 / secs, "unit": "candidates/s", "cores": k,
                                "kind": "reference", "sample": f"{k} pairs at the reference's 1 degree step, {secs:.1f} s",
                                "poses_bit_identical_to_gpu": bool(np.array_equal(rx, g[1]) and np.array_equal(ry, g[2]) and np.array_equal(rt, g[3]))}
        m1.close()
    m.close()
    return line


def bench_solver(cpu):
    """The matcher the Nautilus node itself constructs: CSM(30, 2, 0.3, 0.01) through the
    5-argument overload with a 90 degree window (src/optimization/solver.cc:56, 633-638),
    as a loop-closure batch: 400 scans, every target matched by 5 queries."""
    n_scans, group = 400, 5
    xy, offs, poses = synth.make_scan_set(n_scans, group=group, seed=5, return_poses=True)
    pa, pb = synth.make_pair_list(n_scans, group=group)        # 2000 pairs
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.01)
    rng = np.random.default_rng(5)
    est = poses[:, 2] + rng.normal(0, 0.05, n_scans)           # pose-graph angle estimates
    wins = np.asarray([m.rotation_window(est[a], est[b], np.pi / 2) for a, b in zip(pa, pb)])
    m.upload_scans(xy, offs)
    m.match_batch(pa[:50], pb[:50], wins[:50, 0], wins[:50, 1])
    runs = []
    for _ in range(3):
        t0 = time.perf_counter()
        s, x, y, th, st = m.match_batch(pa, pb, wins[:, 0], wins[:, 1], with_stats=True)
        wall = time.perf_counter() - t0
        runs.append((m.last_timings(), wall))
    tm, wall = sorted(runs, key=lambda r: r[0]["ms_total"])[1]
    cand = int(st[:, 0].sum() + st[:, 1].sum())
    dense = 8.0 * (m.w_high ** 2 + m.w_low ** 2) * n_scans
    line = {"config": "solver matcher CSM(30, 2, 0.3, 0.01), 5-arg overload, 90 deg window "
                      "(solver.cc:56,633-638): 2000 pairs over 400 scans",
            "metric": "candidate_transforms_scored_per_sec", "value": cand / (tm["ms_total"] * 1e-3),
            "pairs_per_sec": len(pa) / (tm["ms_total"] * 1e-3), "ms": tm["ms_total"], "ms_wall": wall * 1e3,
            "kernels_ms": {k: tm[k] for k in ("ms_tables", "ms_low", "ms_refine")},
            "waves": tm["waves"], "w_high": m.w_high, "hi_candidates_per_window": int(st[0, 1] // max(1, st[0, 2])),
            "mean_pops": float(st[:, 2].mean()), "spec_windows": tm["spec_windows"],
            "low_score_kernel_fp64_add_frac": low_fp64_frac(st, offs, pa, tm["ms_low"]),
            "tables": {"dense_layout_GBps": dense / (tm["ms_tables"] * 1e-3) / 1e9,
                       "dense_layout_frac_of_hbm": dense / (tm["ms_tables"] * 1e-3) / 1e9 / PEAK,
                       "note": "dense-layout bytes are the reference's 8 (W^2 + Wl^2) per table; the "
                               "kernels move the cropped rectangle only - actual traffic: "
                               "profiles/traffic.json"}}
    if cpu:
        from oracle import ref
        k = cores()
        rm = ref.RefMatcher(30, 2, 0.3, 0.01)
        idx = np.linspace(0, len(pa) - 1, k).astype(np.int64)
        secs, rs, rx, ry, rt = rm.match_batch(xy, offs, pa[idx], pb[idx], wins[idx, 0], wins[idx, 1],
                                              threads=k)
        c1 = int(st[idx, 0].sum() + st[idx, 1].sum())
        line["cpu_baseline"] = {"value": c1 / secs, "unit": "candidates/s", "pairs_per_sec": k / secs,
                                "cores": k, "kind": "reference",
                                "sample": f"{k} of {len(pa)} pairs, one pair per thread, {secs:.1f} s",
                                "poses_bit_identical_to_gpu": bool(
                                    np.array_equal(rx, x[idx]) and np.array_equal(ry, y[idx])
                                    and np.array_equal(rt, th[idx])),
                                "worst_score_rel_err": float(np.max(np.abs(rs - s[idx]) / np.abs(rs)))}
    m.close()
    return line


def bench_uncertainty(cpu):
    """bag_avg_matcher.cc:104-141's pattern: every base cloud against its previous clouds,
    GetTransformation over the full rotation + the 2x2 GetUncertaintyMatrix at the matched
    rotation (GetLocalUncertaintyStats, cc:375-398), as two batched device passes."""
    n_scans, prev = 200, 10
    xy, offs = synth.make_scan_set(n_scans, group=20, seed=7, trans_range=2.0, max_rot=0.3)
    a_ids, b_ids = [], []
    for s in range(n_scans):
        g0 = (s // 20) * 20
        for q in range(max(g0, s - prev), s):
            a_ids.append(s); b_ids.append(q)
    a_ids, b_ids = np.asarray(a_ids, np.int32), np.asarray(b_ids, np.int32)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m.upload_scans(xy, offs)
    m.match_batch(a_ids[:20], b_ids[:20], 0.0, 2 * np.pi)
    t0 = time.perf_counter()
    _, _, _, th = m.match_batch(a_ids, b_ids, 0.0, 2 * np.pi)
    t1 = time.perf_counter()
    us = m.uncertainty_batch(a_ids, b_ids, th.astype(np.float64))
    t2 = time.perf_counter()
    line = {"config": f"local uncertainty stats, bag_avg_matcher pattern: {n_scans} base clouds x <= {prev} "
                      f"previous clouds = {len(a_ids)} pairs, full rotation + 2x2 covariance",
            "metric": "pairs_per_sec", "value": len(a_ids) / (t2 - t0), "ms_match": 1e3 * (t1 - t0),
            "ms_uncertainty": 1e3 * (t2 - t1), "n_pairs": int(len(a_ids))}
    if cpu:
        from concurrent.futures import ThreadPoolExecutor
        from oracle import ref
        rm = ref.RefMatcher(30, 2, 0.3, 0.03)
        k = cores()
        idx = np.linspace(0, len(a_ids) - 1, k).astype(np.int64)

        def one(i):
            a = xy[offs[a_ids[i]]:offs[a_ids[i] + 1]]
            b = xy[offs[b_ids[i]]:offs[b_ids[i] + 1]]
            r = rm.match(a, b)
            return r[3], rm.uncertainty2(a, b, float(r[3]))
        t0 = time.perf_counter()
        with ThreadPoolExecutor(max_workers=k) as ex:
            res = list(ex.map(one, idx))
        secs = time.perf_counter() - t0
        worst = max(float(np.abs(us[i] - u).max() / np.abs(u).max()) for i, (_, u) in zip(idx, res))
        line["cpu_baseline"] = {"value": k / secs, "unit": "pairs/s", "cores": k, "kind": "reference",
                                "sample": f"{k} pairs, one pair per thread, {secs:.1f} s",
                                "rotations_identical_to_gpu": bool(all(np.float32(r) == th[i] for i, (r, _) in zip(idx, res))),
                                "worst_matrix_rel_err": worst}
    m.close()
    return line


def bench_nn(cpu):
    n_scans, window = 2000, 10
    xy, offs, poses = synth.make_scan_set(n_scans, group=40, seed=6, trans_range=4.0, max_rot=0.5,
                                          return_poses=True)
    rng = np.random.default_rng(3)
    est = poses + np.concatenate([rng.normal(0, 0.05, (n_scans, 2)), rng.normal(0, 0.01, (n_scans, 1))], 1)
    src, tgt = [], []
    for i in range(n_scans):                      # BuildOptimizationOverWindow: 10 previous nodes
        g0 = (i // 40) * 40
        for j in range(max(g0, i - window), i):
            src.append(i); tgt.append(j)
    src, tgt = np.asarray(src, np.int32), np.asarray(tgt, np.int32)
    tr = np.stack([pose_transform(est[s], est[t]) for s, t in zip(src, tgt)])
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m.upload_scans(xy, offs)
    t0 = time.perf_counter()
    m.nn_build(0.25)
    build_s = time.perf_counter() - t0
    m.nn_match_batch(src[:100], tgt[:100], tr[:100], 0.25)
    ms, walls = [], []
    for _ in range(3):
        t0 = time.perf_counter()
        idx, o = m.nn_match_batch(src, tgt, tr, 0.25)
        walls.append(time.perf_counter() - t0)
        ms.append(m.nn_last_stats()["ms_kernel"])
    st = m.nn_last_stats()
    t = float(np.median(ms)) * 1e-3
    pts = (offs[1:] - offs[:-1]).astype(np.int64)
    alg = float((8 * pts[src] + 8 * pts[tgt] + 4 * pts[src]).sum())
    line = {"config": "point-to-point NN correspondences, 2000-pose trajectory slice, window 10 "
                      "(BASELINE configs[4] shape; 20k poses = 10x this batch)",
            "metric": "nn_pairs_per_sec", "value": len(src) / t, "queries_per_sec": len(idx) / t,
            "ms_kernel": t * 1e3, "ms_wall_with_d2h": float(np.median(walls)) * 1e3,
            "n_pairs": int(len(src)), "n_queries": int(len(idx)), "matched_frac": float((idx >= 0).mean()),
            "tree_walk_frac": st["n_walked"] / max(1, len(idx)), "index_build_s": build_s,
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_pair": alg / len(src),
                         "achieved": alg / t / 1e9, "peak": PEAK, "unit": "GB/s",
                         "frac": alg / t / 1e9 / PEAK}}
    if cpu:
        from oracle import ref
        k = min(len(src), 4000)
        secs, ridx, ro = ref.nn_match_batch(xy, offs, src[:k], tgt[:k], tr[:k], 0.25, threads=cores())
        line["cpu_baseline"] = {"value": k / secs, "unit": "pairs/s", "cores": cores(),
                                "kind": "reference", "sample": f"{k} pairs (trees prebuilt, untimed), {secs:.1f} s",
                                "indices_identical_to_gpu": bool(np.array_equal(ridx, idx[:ro[-1]]))}
    m.close()
    return line


def bench_frontend(cpu):
    from nautilus_b200 import synth_ranges as sr
    n = 20000                                        # BASELINE configs[4]: a 20k-pose trajectory
    r = sr.make_ranges(n, seed=1)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    lines = []
    for variant, name in ((0, "nautilus::LaserScanToPointCloud"), (1, "csm LaserScanToPointCloud + FilterGlancing")):
        ms, walls = [], []
        for _ in range(4):
            t0 = time.perf_counter()
            m.upload_ranges(r, sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN, sr.MAX_RANGE,
                            variant=variant)
            walls.append(time.perf_counter() - t0)
            ms.append(m.frontend_last_ms())
        t = float(np.median(ms[1:])) * 1e-3
        n_pts = int(m._scan_sizes.sum())
        # count pass reads the ranges, write pass reads them again and writes the points
        alg = 4.0 * r.size + 8.0 * n_pts
        line = {"config": f"front end: {name}, {n} scans x {r.shape[1]} beams",
                "metric": "scans_per_sec", "value": n / t, "ms_kernels": t * 1e3,
                "ms_wall_with_h2d": float(np.median(walls[1:])) * 1e3, "points_out": n_pts,
                "roofline": {"bound": "hbm", "algorithmic_bytes_per_scan": alg / n,
                             "achieved": alg / t / 1e9, "peak": PEAK, "unit": "GB/s",
                             "frac": alg / t / 1e9 / PEAK}}
        if cpu and variant == 1:
            from oracle import ref
            k = 2000
            t0 = time.perf_counter()
            tot = 0
            for s in range(k):
                tot += len(ref.csm_laser_to_points(r[s], sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC,
                                                   sr.RANGE_MIN, sr.MAX_RANGE, False))
            secs = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": k / secs, "unit": "scans/s", "cores": 1, "kind": "reference",
                                    "sample": f"{k} scans, one thread, {secs:.2f} s",
                                    "counts_identical_to_gpu": bool(tot == int(m._scan_sizes[:k].sum()))}
        lines.append(line)
    t0 = time.perf_counter()
    sc = m.scatter_scores()
    lines.append({"config": f"LC candidate filter: scatter score of {n} scans", "metric": "scans_per_sec",
                  "value": n / (time.perf_counter() - t0), "note": "wall time incl. D2H of the scores",
                  "frac_above_0.70": float((sc >= 0.70).mean())})
    m.close()
    return lines


def bench_residuals(cpu):
    from nautilus_b200.residuals import evaluate_residual_blocks, last_kernel_ms
    rng = np.random.default_rng(0)
    nb, per = 20000, 800                              # 1.6e7 correspondences
    offs = np.arange(nb + 1, dtype=np.int64) * per
    sp = rng.uniform(-20, 20, size=(nb, 3))
    tp = sp + rng.normal(0, 0.2, size=(nb, 3))
    p = rng.uniform(-25, 25, size=(nb * per, 2)).astype(np.float32)
    q = (p + rng.normal(0, 0.1, size=p.shape)).astype(np.float32)
    a = rng.uniform(-np.pi, np.pi, size=(2, nb * per))
    ns = np.stack([np.cos(a[0]), np.sin(a[0])], 1).astype(np.float32)
    nt = np.stack([np.cos(a[1]), np.sin(a[1])], 1).astype(np.float32)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    lines = []
    for kind, name in ((0, "LIDARPointResidual"), (1, "LIDARNormalResidual")):
        ms, walls = [], []
        for _ in range(3):
            t0 = time.perf_counter()
            evaluate_residual_blocks(m, kind, sp, tp, offs, p, q, ns, nt)
            walls.append(time.perf_counter() - t0)
            ms.append(last_kernel_ms(m))
        t = float(np.median(ms)) * 1e-3
        n = nb * per
        alg = n * ((16 if kind == 0 else 32) + 16 + 96.0)
        line = {"config": f"residual blocks: {name}, {nb} blocks x {per} correspondences, residuals + both Jacobians",
                "metric": "correspondences_per_sec", "value": n / t, "ms_kernel": t * 1e3,
                "ms_wall_with_copies": float(np.median(walls)) * 1e3,
                "roofline": {"bound": "hbm", "algorithmic_bytes_per_correspondence": alg / n,
                             "achieved": alg / t / 1e9, "peak": PEAK, "unit": "GB/s",
                             "frac": alg / t / 1e9 / PEAK}}
        if cpu:
            from oracle import residual_port as rp
            k = 200
            t0 = time.perf_counter()
            rp.evaluate(kind, sp[:k], tp[:k], offs[:k + 1], p[:k * per], q[:k * per], ns[:k * per], nt[:k * per])
            secs = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": k * per / secs, "unit": "correspondences/s", "cores": 1,
                                    "kind": "port", "sample": f"{k} blocks, numpy dual numbers, {secs:.2f} s"}
        lines.append(line)
    m.close()
    return lines


def main():
    which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["tables", "wide", "solver", "uncertainty", "nn", "frontend", "residuals"]
    cpu = "--cpu" in sys.argv
    for w in which:
        out = {"tables": bench_tables, "wide": bench_wide, "nn": bench_nn, "solver": bench_solver, "uncertainty": bench_uncertainty,
               "frontend": bench_frontend, "residuals": bench_residuals}[w](cpu)
        for line in (out if isinstance(out, list) else [out]):
            print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
