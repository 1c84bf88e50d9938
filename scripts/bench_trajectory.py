"""BASELINE configs[4]: synthetic 20k-pose multi-robot trajectory (4 robots x 5000
poses on a shared 60 x 40 m plan, 0.25 m / <=10 degrees between poses, default
parameters as in ut_multirobot_sim_slam_config.lua) through the device pipeline:

  ranges --K9--> clouds --scatter score + GetLCCandidates--> keyframes
         --all-candidate CSM (90 degree restriction, solver.cc:638)--> (pose, score)
  clouds --K7, growing window 1..10 (55 pair matchings per node)--> correspondences
         --K10--> residuals + Jacobians of the widest window round

One JSON line: per-stage wall times of the C-ABI calls (host buffers, copies
included) and the kernel times the library reports.  `--cpu` times the compiled
reference (oracle/_ref) on bounded samples of the CSM and NN stages.

  python scripts/bench_trajectory.py [--poses 5000] [--cpu]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from nautilus_b200 import CorrelativeScanMatcher, synth_ranges as sr  # noqa: E402
from nautilus_b200.residuals import (evaluate_residual_blocks, evaluate_resident_blocks,  # noqa: E402
                                     last_kernel_ms)

HALF_W, HALF_H = 30.0, 20.0


def make_plan(seed=0):
    rng = np.random.default_rng(seed)
    segs = [(-HALF_W, -HALF_H, HALF_W, -HALF_H), (HALF_W, -HALF_H, HALF_W, HALF_H),
            (HALF_W, HALF_H, -HALF_W, HALF_H), (-HALF_W, HALF_H, -HALF_W, -HALF_H)]
    for _ in range(36):
        x, y = rng.uniform(-HALF_W + 2, HALF_W - 2), rng.uniform(-HALF_H + 2, HALF_H - 2)
        ln = rng.uniform(3.0, 12.0)
        if rng.random() < 0.5:
            segs.append((x, y, min(x + ln, HALF_W), y))
        else:
            segs.append((x, y, x, min(y + ln, HALF_H)))
    return np.asarray(segs, np.float64)


def make_trajectories(n_robots, n_poses, seed=1):
    """0.25 m steps (translation_change_for_lidar), heading change <= 10 degrees
    (rotation_change_for_lidar), turning away from the outer walls."""
    rng = np.random.default_rng(seed)
    out = []
    for r in range(n_robots):
        p = np.array([rng.uniform(-HALF_W + 3, HALF_W - 3), rng.uniform(-HALF_H + 3, HALF_H - 3)])
        th = rng.uniform(-np.pi, np.pi)
        poses = np.empty((n_poses, 3))
        for k in range(n_poses):
            poses[k] = (p[0], p[1], th)
            th += rng.uniform(-1, 1) * np.pi / 18
            nxt = p + 0.25 * np.array([np.cos(th), np.sin(th)])
            if abs(nxt[0]) > HALF_W - 1.5 or abs(nxt[1]) > HALF_H - 1.5:
                th += np.pi / 18 * 3 * (1 if rng.random() < 0.5 else -1)  # keep turning
                nxt = p + 0.1 * np.array([np.cos(th), np.sin(th)])
                nxt = np.clip(nxt, [-HALF_W + 1.0, -HALF_H + 1.0], [HALF_W - 1.0, HALF_H - 1.0])
            p = nxt
        out.append(poses)
    return np.concatenate(out, axis=0)


def cast_ranges(plan, poses, seed=2, chunk=256):
    rng = np.random.default_rng(seed)
    n = len(poses)
    out = np.empty((n, sr.N_BEAMS), np.float32)
    beam = float(sr.ANGLE_MIN) + np.arange(sr.N_BEAMS) * float(sr.ANGLE_INC)
    ax, ay = plan[:, 0], plan[:, 1]
    ex, ey = plan[:, 2] - ax, plan[:, 3] - ay
    for s in range(0, n, chunk):
        ps = poses[s:s + chunk]
        ang = ps[:, 2:3] + beam[None, :]
        dx, dy = np.cos(ang)[..., None], np.sin(ang)[..., None]
        rx, ry = ax[None, None, :] - ps[:, 0, None, None], ay[None, None, :] - ps[:, 1, None, None]
        den = dx * ey - dy * ex
        with np.errstate(divide="ignore", invalid="ignore"):
            t = (rx * ey - ry * ex) / den
            u = (rx * dy - ry * dx) / den
        hit = (np.abs(den) > 1e-12) & (t > 1e-3) & (u >= 0) & (u <= 1)
        r = np.where(hit, t, np.inf).min(axis=2)
        out[s:s + chunk] = (r + rng.normal(0, 0.01, r.shape)).astype(np.float32)
    return out


def transforms(est, src, tgt):
    """target^-1 * source as row-major 2x3 float (solver.cc:148-159), vectorised."""
    s, t = est[src], est[tgt]
    dth = s[:, 2] - t[:, 2]
    c, sn = np.cos(dth), np.sin(dth)
    ct, st = np.cos(t[:, 2]), np.sin(t[:, 2])
    dx, dy = s[:, 0] - t[:, 0], s[:, 1] - t[:, 1]
    m = np.stack([c, -sn, ct * dx + st * dy, sn, c, -st * dx + ct * dy], axis=1)
    return np.ascontiguousarray(m, np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--poses", type=int, default=5000, help="poses per robot")
    ap.add_argument("--robots", type=int, default=4)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--stage-cells", type=int, default=0, help="tuning: staged low-res table budget")
    ap.add_argument("--only-csm", action="store_true")
    ap.add_argument("--high-res", type=float, default=0.01,
                    help="hi-res cell of the all-candidate CSM stage; 0.01 is what the solver "
                         "constructs (solver.cc:56,633); 0.03 is the CSM drivers' value")
    args = ap.parse_args()
    plan = make_plan()
    poses = make_trajectories(args.robots, args.poses)
    n = len(poses)
    rng = np.random.default_rng(5)
    est = poses + np.concatenate([rng.normal(0, 0.05, (n, 2)), rng.normal(0, 0.01, (n, 1))], 1)
    ranges = cast_ranges(plan, poses)
    robot = np.arange(n) // args.poses
    line = {"config": f"{args.robots} robots x {args.poses} poses on a 60 x 40 m plan "
                      "(BASELINE configs[4]), 1080 beams, default_config.lua parameters",
            "n_poses": n, "stages": {}}
    st = line["stages"]
    m = CorrelativeScanMatcher(30, 2, 0.3, args.high_res, stage_cells=args.stage_cells)
    line["csm_matcher"] = f"CSM(30, 2, 0.3, {args.high_res}), 5-argument overload, 90 degree window"
    m.upload_ranges(ranges[:64], sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN, sr.MAX_RANGE)

    t0 = time.perf_counter()
    m.upload_ranges(ranges, sr.ANGLE_MIN, sr.ANGLE_MAX, sr.ANGLE_INC, sr.RANGE_MIN, sr.MAX_RANGE)
    st["ranges_to_clouds"] = {"ms_wall": 1e3 * (time.perf_counter() - t0), "ms_kernels": m.frontend_last_ms(),
                              "points": int(m._scan_sizes.sum())}
    t0 = time.perf_counter()
    keys = m.GetLCCandidates(est[:, :2])
    st["lc_candidate_filter"] = {"ms_wall": 1e3 * (time.perf_counter() - t0), "keyframes": int(len(keys))}

    # all-candidate CSM: keyframe pairs within lc_base_max_range + lc_max_range_scaling * index gap
    kp = est[keys, :2]
    d = np.hypot(kp[:, None, 0] - kp[None, :, 0], kp[:, None, 1] - kp[None, :, 1])
    gap = np.abs(keys[:, None] - keys[None, :])
    ii, jj = np.nonzero((d <= 3.5 + 0.01 * gap) & (keys[:, None] < keys[None, :]))
    a_ids, b_ids = keys[jj].astype(np.int32), keys[ii].astype(np.int32)
    win = np.array([m.rotation_window(est[a, 2], est[b, 2], np.pi / 2) for a, b in zip(a_ids, b_ids)],
                   np.float64).reshape(-1, 2)
    # the NN index is not needed by the CSM stage: its host half (the reference's KD-tree
    # layout) runs on library threads underneath it
    t_begin = time.perf_counter()
    if not args.only_csm:
        m.nn_build_begin(0.25)
    if len(a_ids):
        m.match_batch(a_ids[:8], b_ids[:8], win[:8, 0], win[:8, 1])
        t0 = time.perf_counter()
        res = m.match_batch(a_ids, b_ids, win[:, 0], win[:, 1], with_stats=True)
        wall = time.perf_counter() - t0
        tm = m.last_timings()
        cand = int(res[4][:, 0].sum() + res[4][:, 1].sum())
        st["csm_all_candidates"] = {"pairs": int(len(a_ids)), "ms_wall": 1e3 * wall, "ms_device": tm["ms_total"],
                                    "candidates_per_sec": cand / (tm["ms_total"] * 1e-3),
                                    "pairs_per_sec": len(a_ids) / (tm["ms_total"] * 1e-3),
                                    "ms_tables": tm["ms_tables"], "ms_low": tm["ms_low"], "ms_refine": tm["ms_refine"],
                                    "speculative_windows": tm["spec_windows"], "mean_pops": float(res[4][:, 2].mean()),
                                    "max_pops": int(res[4][:, 2].max())}
    else:
        st["csm_all_candidates"] = {"pairs": 0}

    if args.only_csm:
        print(json.dumps(line), flush=True)
        return
    # growing window 1..10 (OptimizeOverGrowingWindow): sum_w w = 55 pair matchings per node
    t0 = time.perf_counter()
    m.nn_build(0.25)
    st["nn_index_build"] = {"ms_wall_after_csm": 1e3 * (time.perf_counter() - t0),
                            "ms_since_begin": 1e3 * (time.perf_counter() - t_begin),
                            "note": "host KD-tree layout started before the CSM stage "
                                    "(nautilus_nn_build_begin) and joined here"}
    tot_pairs = tot_q = 0
    ms_k = ms_w = 0.0
    last = None
    for w in range(1, 11):
        src = np.repeat(np.arange(n), w)
        tgt = src - np.tile(np.arange(1, w + 1), n)
        ok = (tgt >= 0) & (robot[np.maximum(tgt, 0)] == robot[src])
        src, tgt = src[ok].astype(np.int32), tgt[ok].astype(np.int32)
        tr = transforms(est, src, tgt)
        for c0 in range(0, len(src), 60000):     # bounded result buffers
            sl = slice(c0, c0 + 60000)
            t0 = time.perf_counter()
            idx, offs = m.nn_match_batch(src[sl], tgt[sl], tr[sl], 0.25)[:2]
            ms_w += 1e3 * (time.perf_counter() - t0)
            ms_k += m.nn_last_stats()["ms_kernel"]
            tot_pairs += len(src[sl]); tot_q += len(idx)
            if w == 10 and c0 == 0:
                last = (src[sl], tgt[sl], idx, offs, tr[sl])
    st["nn_growing_window"] = {"pair_matchings": int(tot_pairs), "queries": int(tot_q), "ms_kernel": ms_k,
                               "ms_wall_with_copies": ms_w, "pairs_per_sec_kernel": tot_pairs / (ms_k * 1e-3),
                               "queries_per_sec_kernel": tot_q / (ms_k * 1e-3)}

    # residual blocks of one 60k-pair slice of the widest round (point residuals)
    src, tgt, idx, offs, tr_last = last
    xy, soffs = m.download_scans()
    t0 = time.perf_counter()
    hit = idx >= 0
    blk = np.searchsorted(offs, np.arange(len(idx)), side="right") - 1
    sp = xy[soffs[src[blk[hit]]] + (np.arange(len(idx)) - offs[blk])[hit]]
    tp = xy[soffs[tgt[blk[hit]]] + idx[hit]]
    roffs = np.zeros(len(src) + 1, np.int64)
    np.cumsum(np.bincount(blk[hit], minlength=len(src)), out=roffs[1:])
    host_ms = 1e3 * (time.perf_counter() - t0)
    t0 = time.perf_counter()
    res, js, jt = evaluate_residual_blocks(m, 0, est[src], est[tgt], roffs, sp, tp)
    st["residual_blocks"] = {"blocks": int(len(src)), "correspondences": int(len(sp)),
                             "ms_host_gather": host_ms, "ms_wall_with_copies": 1e3 * (time.perf_counter() - t0),
                             "ms_kernel": last_kernel_ms(m), "rms_residual_m": float(np.sqrt((res ** 2).mean()))}

    # the same slice with the correspondences kept on the device (nn_match_resident ->
    # evaluate_resident_blocks): only counts, poses and residuals / Jacobians cross PCIe
    t0 = time.perf_counter()
    co = m.nn_match_resident(src, tgt, tr_last, 0.25)
    ms_match = 1e3 * (time.perf_counter() - t0)
    ms_match_k = m.nn_last_stats()["ms_kernel"]
    t0 = time.perf_counter()
    r_only = evaluate_resident_blocks(m, 0, est[src], est[tgt], co, jacobians=False)
    ms_res = 1e3 * (time.perf_counter() - t0)
    t0 = time.perf_counter()
    r_all = evaluate_resident_blocks(m, 0, est[src], est[tgt], co)
    ms_all = 1e3 * (time.perf_counter() - t0)
    t0 = time.perf_counter()
    r_pin = evaluate_resident_blocks(m, 0, est[src], est[tgt], co, pinned=True)   # allocates the pinned buffers
    ms_pin_first = 1e3 * (time.perf_counter() - t0)
    t0 = time.perf_counter()
    r_pin = evaluate_resident_blocks(m, 0, est[src], est[tgt], co, out=r_pin)     # a later solver iteration
    ms_pin = 1e3 * (time.perf_counter() - t0)
    st["resident_chain"] = {"blocks": int(len(src)), "correspondences": int(co[-1]),
                            "ms_with_jacobians_pinned_first_wall": ms_pin_first,
                            "ms_with_jacobians_pinned_wall": ms_pin,
                            "pinned_identical": bool(np.array_equal(r_pin[0], res) and np.array_equal(r_pin[1], js)
                                                     and np.array_equal(r_pin[2], jt)),
                            "ms_match_wall": ms_match, "ms_match_kernel": ms_match_k,
                            "ms_residuals_only_wall": ms_res, "ms_with_jacobians_wall": ms_all,
                            "ms_kernel": last_kernel_ms(m),
                            "identical_to_host_chain": bool(np.array_equal(co, roffs) and np.array_equal(r_only, res)
                                                            and np.array_equal(r_all[1], js)
                                                            and np.array_equal(r_all[2], jt))}

    if args.cpu:
        from oracle import ref
        cores = len(os.sched_getaffinity(0))
        cpu = {"cores": cores, "kind": "reference"}
        if len(a_ids):
            k = min(len(a_ids), 4 * cores)
            k = min(k, 2 * cores) if args.high_res < 0.02 else k   # ~3 s per pair at 0.01 m
            secs, rs, rx, ry, rt = ref.RefMatcher(30, 2, 0.3, args.high_res).match_batch(
                xy, soffs, a_ids[:k], b_ids[:k], win[:k, 0].copy(), win[:k, 1].copy(), threads=cores)
            g = m.match_batch(a_ids[:k], b_ids[:k], win[:k, 0], win[:k, 1])
            cpu["csm"] = {"pairs_per_sec": k / secs, "sample": f"{k} pairs, {secs:.1f} s",
                          "poses_bit_identical_to_gpu": bool(np.array_equal(rx, g[1]) and np.array_equal(ry, g[2])
                                                             and np.array_equal(rt, g[3]))}
        k = 4000
        secs, ridx, ro = ref.nn_match_batch(xy, soffs, src[:k], tgt[:k], transforms(est, src[:k], tgt[:k]),
                                            0.25, threads=cores)
        cpu["nn"] = {"pairs_per_sec": k / secs, "sample": f"{k} pairs (trees prebuilt, untimed), {secs:.2f} s",
                     "indices_identical_to_gpu": bool(np.array_equal(ridx, idx[:ro[-1]]))}
        line["cpu_baseline"] = cpu
    m.close()
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
