"""bench.py -- candidate transforms scored per second for the loop-closure batch
(BASELINE.json configs[2]: 10k synthetic scan pairs, multi-res branch search).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch: build the lookup tables of
the batch's target scans and match its (query, target) pairs: 10 000 pairs over
2 000 scans per GPU.  With N GPUs there is ONE batch of N x 10 000 pairs which the
library partitions over the ranks (nautilus_csm_match_batch_sharded: sort by target,
balanced contiguous blocks, one ncclAllGather of the 48-byte records); the
strong-scaling split of a single 10 000-pair batch is measured in the same run.

One JSON line is printed by rank 0; see DESIGN.md "Measurement" for every key.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PARAMS = (30.0, 2.0, 0.3, 0.03)   # CSM(30, 2, 0.3, 0.03): bag_pair_matcher.cc:99-100
ROT_WINDOW = (-0.5, 0.5)          # default_config CSM window of BASELINE configs[0]
N_SCANS, GROUP = 2000, 5          # 2000 scans, every target matched by 5 queries
METRIC = "candidate_transforms_scored_per_sec"
UNIT = "candidates/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scans", type=int, default=N_SCANS)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N > 1: weak = one batch of N x 10 000 pairs partitioned over the N GPUs "
                         "(headline); strong = the 10 000-pair batch split N ways.  The other split "
                         "is measured too and reported under 'other_split'")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--stage-cells", type=int, default=0, help="tuning: staged low-res table budget")
    return ap.parse_args()


def workload_config(n_scans, n_gpus):
    return {
        "workload": "loop-closure candidate batch (BASELINE configs[2])",
        "pairs_per_gpu": n_scans * GROUP, "scans_per_gpu": n_scans,
        "beams": 1080, "matcher": "CSM(range=30, trans=2, low=0.3, high=0.03)",
        "rotation_window_rad": list(ROT_WINDOW), "rot_step_deg": 1.0,
        "parallelism": "single GPU",
        "l2": "per-step working set (tables) ~10 GB >> 126 MB L2, no flush needed",
    }


def make_batch(n_scans, seed):
    from nautilus_b200 import synth
    xy, offs = synth.make_scan_set(n_scans, group=GROUP, seed=seed)
    pa, pb = synth.make_pair_list(n_scans, group=GROUP)
    return xy, offs, pa, pb


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (NVML in a
    thread, every 20 ms; NVML is initialised before the warm-up so that its start-up
    does not fall into the timed region).  Falls back to one nvidia-smi query."""
    REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
               "sw_power_cap": 0x4}

    def __init__(self, index):
        self.index, self.samples, self.run, self.t, self.h = index, [], False, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None

    def _loop(self):
        nv = self.nv
        while self.run:
            try:
                self.samples.append((float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)),
                                     int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                                     if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons")
                                     else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))))
            except Exception:
                pass
            time.sleep(0.02)

    def start(self):
        if self.h is None:
            return
        self.samples, self.run = [], True
        self.t = threading.Thread(target=self._loop, daemon=True)
        self.t.start()

    def stop(self):
        if self.h is None:
            return self._smi_once()
        self.run = False
        self.t.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_sm, "reasons": ["no samples"]}
        sm = sorted(v for v, _ in self.samples)
        bits = 0
        for _, r in self.samples:
            bits |= r
        reasons = sorted(k for k, b in self.REASONS.items() if bits & b)
        busy = sm[len(sm) // 2:]  # upper half = samples under load
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": self.max_sm, "reasons": reasons,
                "samples": len(sm), "source": "nvml"}

    def _smi_once(self):
        try:
            out = subprocess.run(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm",
                 "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=20).stdout
            f = [float(v) for v in out.strip().split(",")]
            return {"sm_mhz": f[0], "sm_max_mhz": f[1], "reasons": [], "samples": 1,
                    "source": "nvidia-smi after the timed region (NVML unavailable)"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}


def ncu_traffic():
    """Per-kernel DRAM traffic of the committed ncu capture (profiles/traffic.json,
    written by scripts/ncu_summary.py); None if absent."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(p))
    except Exception:
        return None


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


# ---- CPU side: the reference's own code (oracle/_ref) or the oracle port -------
def count_candidates(xy, offs, pa, pb, idx, threads):
    """Candidates the reference scores for the sampled pairs (low + high), counted
    by the instrumented oracle port on host threads; NOT timed."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import port
    pm = port.PortMatcher(*PARAMS)

    def one(k):
        a = xy[offs[pa[k]]:offs[pa[k] + 1]]
        b = xy[offs[pb[k]]:offs[pb[k] + 1]]
        _, st = pm.match(a, b, *ROT_WINDOW, with_stats=True)
        return int(st[0] + st[1])

    with ThreadPoolExecutor(max_workers=threads) as ex:
        return list(ex.map(one, idx))


def cpu_reference_run(xy, offs, pa, pb, idx, threads):
    """Times GetTransformation over the sampled pairs with one pair per host thread
    (the pattern of bag_avg_matcher.cc:104).  Returns (seconds, kind, results)."""
    from oracle import build, ref
    rmin = np.full(len(idx), ROT_WINDOW[0])
    rmax = np.full(len(idx), ROT_WINDOW[1])
    if build.build_ref() is not None:
        rm = ref.RefMatcher(*PARAMS)
        secs, score, x, y, th = rm.match_batch(xy, offs, pa[idx], pb[idx], rmin, rmax,
                                               threads=threads)
        return secs, "reference", (score, x, y, th)
    # fallback: the plain-C restatement on a thread pool
    from concurrent.futures import ThreadPoolExecutor
    from oracle import port
    pm = port.PortMatcher(*PARAMS)

    def one(k):
        return pm.match(xy[offs[pa[k]]:offs[pa[k] + 1]], xy[offs[pb[k]]:offs[pb[k] + 1]],
                        *ROT_WINDOW)
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        res = list(ex.map(one, idx))
    secs = time.perf_counter() - t0
    r = np.asarray(res, np.float64)
    return secs, "port", (r[:, 0], r[:, 1].astype(np.float32), r[:, 2].astype(np.float32),
                          r[:, 3].astype(np.float32))


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference_arm(args):
    """--impl reference: the reference's CPU CSM on the box's host cores, all
    threads, each step a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    xy, offs, pa, pb = make_batch(min(args.scans, 200), seed=0)
    n_sample = min(len(pa), 8 * cores)
    idx = np.linspace(0, len(pa) - 1, n_sample).astype(np.int64)
    cands = count_candidates(xy, offs, pa, pb, idx, cores)
    kind = "reference"
    for _ in range(args.warmup):
        cpu_reference_run(xy, offs, pa, pb, idx[:max(1, cores // 2)], cores)
    total = 0.0
    for _ in range(args.steps):
        secs, kind, _ = cpu_reference_run(xy, offs, pa, pb, idx, cores)
        total += secs
    value = sum(cands) * args.steps / total
    sample = (f"{n_sample} pairs per step, evenly spaced in a {len(pa)}-pair batch over "
              f"{min(args.scans, 200)} scans of the same generator (the full workload has "
              f"{args.scans * GROUP} pairs; candidates/s extrapolates linearly), one pair per thread, "
              f"tables rebuilt per pair as GetTransformation does")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.scans, args.gpus),
        "pairs_per_sec": n_sample * args.steps / total,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---- the B200 arm --------------------------------------------------------------
def global_batch(args, rank, world, torch, dist, dev, strong):
    """The batch the world works on.  weak: ONE batch of world x 10 000 pairs over
    world x 2 000 scans (every rank synthesises its 2 000 scans, the clouds are
    exchanged once during set-up so that every rank holds the whole batch, as the
    sharded C-ABI call expects); strong: the 10 000-pair batch of the N = 1 run."""
    xy, offs, pa, pb = make_batch(args.scans, seed=0 if strong else rank)
    if world == 1 or strong:
        return xy, offs, pa, pb
    n_pts = torch.tensor([len(xy)], device=dev, dtype=torch.int64)
    all_n = [torch.zeros_like(n_pts) for _ in range(world)]
    dist.all_gather(all_n, n_pts)
    all_n = [int(v) for v in all_n]
    mx = max(all_n)
    buf = torch.zeros((mx, 2), device=dev, dtype=torch.float32)
    buf[:len(xy)] = torch.from_numpy(xy).to(dev)
    sizes = torch.zeros(args.scans, device=dev, dtype=torch.int64)
    sizes[:] = torch.from_numpy(np.diff(offs)).to(dev)
    xy_all = [torch.zeros_like(buf) for _ in range(world)]
    sz_all = [torch.zeros_like(sizes) for _ in range(world)]
    dist.all_gather(xy_all, buf)
    dist.all_gather(sz_all, sizes)
    xy_g = np.concatenate([t[:n].cpu().numpy() for t, n in zip(xy_all, all_n)])
    offs_g = np.zeros(world * args.scans + 1, np.int64)
    offs_g[1:] = np.cumsum(np.concatenate([t.cpu().numpy() for t in sz_all]))
    pa_g = np.concatenate([pa + r * args.scans for r in range(world)]).astype(np.int32)
    pb_g = np.concatenate([pb + r * args.scans for r in range(world)]).astype(np.int32)
    return np.ascontiguousarray(xy_g), offs_g, pa_g, pb_g


def run_ours(args):
    import torch
    import torch.distributed as dist
    from nautilus_b200 import CorrelativeScanMatcher, distributed as D

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a B200: there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    m = CorrelativeScanMatcher(*PARAMS, device=local, stage_cells=args.stage_cells,
                               world_rank=rank, world_size=world)
    stream = torch.cuda.current_stream()
    m.set_stream(stream.cuda_stream)
    if world > 1:
        # the library's own communicator: rank 0 makes the id, torch only carries the bytes
        box = [m.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        m.comm_init(box[0])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    class Leg:
        """One workload (weak or strong batch) with its two step functions."""

        def __init__(self, strong):
            self.strong = strong
            self.xy, self.offs, self.pa, self.pb = global_batch(args, rank, world, torch, dist, dev, strong)
            self.n_pairs = len(self.pa)
            self.rmin = np.full(self.n_pairs, ROT_WINDOW[0])
            self.rmax = np.full(self.n_pairs, ROT_WINDOW[1])
            self.xy_pin = torch.from_numpy(self.xy).pin_memory()
            self.xy_host = self.xy_pin.numpy()
            self.upload_ms = 0.0
            self.h2d = 0

        def make_resident(self):
            m.upload_scans(self.xy_host, self.offs)

        def step_resident(self):
            """Scans resident on every rank; one sharded call for the whole batch."""
            if world == 1:
                return m.match_batch(self.pa, self.pb, self.rmin, self.rmax, with_stats=True)
            return m.match_batch_sharded(self.pa, self.pb, self.rmin, self.rmax, with_stats=True)

        def step_e2e(self):
            """Host buffers: this rank partitions the batch, uploads the scans its block
            touches, matches the block and gathers (nautilus_csm_partition ->
            upload_scans -> match_batch -> nautilus_csm_gather)."""
            if world == 1:
                tu = time.perf_counter()
                m.upload_scans(self.xy_host, self.offs)
                self.upload_ms = round(1e3 * (time.perf_counter() - tu), 3)
                self.h2d = int(self.xy.nbytes + self.offs.nbytes + self.pa.nbytes + self.pb.nbytes
                               + 16 * self.n_pairs)
                return m.match_batch(self.pa, self.pb, self.rmin, self.rmax, with_stats=True)
            w = m.pair_weights(self.pa, self.rmin, self.rmax, self.offs)
            order, bounds = D.partition_pairs(self.pb, w, world)
            mine = order[bounds[rank]:bounds[rank + 1]]
            a, b = self.pa[mine], self.pb[mine]
            s0 = int(min(a.min(), b.min())) if len(mine) else 0
            s1 = int(max(a.max(), b.max())) + 1 if len(mine) else 0
            tu = time.perf_counter()
            sub_offs = self.offs[s0:s1 + 1] - self.offs[s0]
            m.upload_scans(self.xy_host[self.offs[s0]:self.offs[s1]], sub_offs)
            self.upload_ms = round(1e3 * (time.perf_counter() - tu), 3)
            self.h2d = int(8 * (self.offs[s1] - self.offs[s0]) + sub_offs.nbytes + 24 * len(mine))
            self.local = m.match_batch(a - s0, b - s0, self.rmin[mine], self.rmax[mine], with_stats=True)
            self.mine = mine
            return m.gather(order, bounds, with_stats=True)

    def timed(leg, upload, steps):
        """K steps between CUDA events on the launching stream; max over ranks."""
        acc = {"tables": 0.0, "low": 0.0, "refine": 0.0, "gather": 0.0, "launches": 0, "cand": 0,
               "host_plan": 0.0, "host_wait": 0.0, "host_collect": 0.0,
               "step_wall_ms": [], "gpu_span_ms": [], "upload_wall_ms": []}
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(steps):
            ts = time.perf_counter()
            score, x, y, th, st = leg.step_e2e() if upload else leg.step_resident()
            acc["step_wall_ms"].append(round(1e3 * (time.perf_counter() - ts), 3))
            t = m.last_timings()
            acc["gpu_span_ms"].append(round(t["ms_total"], 3))
            if upload:
                acc["upload_wall_ms"].append(leg.upload_ms)
            acc["tables"] += t["ms_tables"]; acc["low"] += t["ms_low"]; acc["refine"] += t["ms_refine"]
            acc["gather"] += t["ms_gather"]
            acc["host_plan"] += t["ms_host_plan"]; acc["host_wait"] += t["ms_host_wait"]
            acc["host_collect"] += t["ms_host_collect"]
            acc["launches"] += t["launches"] + (1 if world > 1 else 0)
            # every rank holds every result: the whole job's candidates, counted once
            acc["cand"] += int(st[:, 0].sum() + st[:, 1].sum())
        e1.record(stream)
        barrier()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        ms = e0.elapsed_time(e1)
        if world > 1:
            v = torch.tensor([ms, wall_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(v, op=dist.ReduceOp.MAX)
            ms, wall_ms = float(v[0]), float(v[1])
        return ms, wall_ms, acc, (score, x, y, th, st)

    def check_gather(leg):
        """Before anything is timed: the gathered records of THIS rank's block are the
        records this rank computed, every rank holds the same gathered result, and the
        sharded call returns it too."""
        if world == 1:
            return {"checked": False}
        full = leg.step_e2e()
        for u, v in zip(full, leg.local):
            assert np.array_equal(u[leg.mine], v), "gathered block differs from the local records"
        leg.make_resident()
        sharded = leg.step_resident()
        for u, v in zip(full, sharded):
            assert np.array_equal(u, v), "sharded call differs from partition + match + gather"
        digest = np.frombuffer(np.ascontiguousarray(full[0]).tobytes(), np.uint8).astype(np.int64)
        h = torch.tensor([int(digest.sum()), int((digest * (np.arange(len(digest)) % 251 + 1)).sum())],
                         device=dev, dtype=torch.int64)
        lo, hi = h.clone(), h.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert bool((lo == hi).all()), "ranks hold different gathered results"
        return {"checked": True, "pairs": int(leg.n_pairs), "block_pairs": int(len(leg.mine))}

    def run_leg(strong, sample_clocks):
        leg = Leg(strong)
        gather_check = check_gather(leg)
        leg.make_resident()
        # NVML start-up, the sampler thread's start and its first queries (which can
        # stall the driver for ~0.2 s) happen during the warm-up; only samples taken
        # inside the timed region are kept
        sampler = ClockSampler(local) if sample_clocks else None
        if sampler is not None and rank == 0:
            sampler.start()
        for _ in range(args.warmup):
            leg.step_resident()
        if sampler is not None:
            sampler.samples = []
        ms, _, acc, last = timed(leg, False, args.steps)          # inputs resident in HBM
        clocks = sampler.stop() if (sampler is not None and rank == 0) else None
        leg.step_e2e()
        ms_e2e_dev, wall_e2e, acc_e2e, _ = timed(leg, True, args.steps)  # host buffers, copies inside
        return leg, ms, acc, last, clocks, ms_e2e_dev, wall_e2e, acc_e2e, gather_check

    main_strong = args.scaling == "strong" and world > 1
    leg, ms, acc, last, clocks, ms_e2e_dev, wall_e2e, acc_e2e, gather_check = run_leg(main_strong, True)
    ms_e2e = max(ms_e2e_dev, wall_e2e)  # the call is synchronous: host time counts
    other = None
    if world > 1:  # the other split of the same machine, as a second line inside this one
        o_leg, o_ms, o_acc, _, _, o_dev, o_wall, o_acc_e2e, o_check = run_leg(not main_strong, False)
        o_e2e = max(o_dev, o_wall)
        other = {"scaling": "weak" if main_strong else "strong",
                 "workload": ("one batch of %d pairs split over %d GPUs" % (o_leg.n_pairs, world)),
                 "value": o_acc["cand"] / (o_ms * 1e-3), "unit": UNIT, "ms_per_step": o_ms / args.steps,
                 "pairs_per_sec": o_leg.n_pairs * args.steps / (o_ms * 1e-3),
                 "e2e": {"value": o_acc_e2e["cand"] / (o_e2e * 1e-3), "unit": UNIT,
                         "ms_per_step": o_e2e / args.steps, "h2d_bytes_per_step": o_leg.h2d,
                         "d2h_bytes_per_step": int(48 * o_leg.n_pairs)},
                 "kernels_ms_per_step_rank0": {k: o_acc[k] / args.steps for k in ("tables", "low", "refine", "gather")},
                 "host_ms_per_step_rank0": {k: o_acc["host_" + k] / args.steps for k in ("plan", "wait", "collect")},
                 "gather_check": o_check}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    score, x, y, th, st = last
    xy, offs, pa, pb, n_pairs = leg.xy, leg.offs, leg.pa, leg.pb, leg.n_pairs
    value = acc["cand"] / (ms * 1e-3)
    e2e_value = acc_e2e["cand"] / (ms_e2e * 1e-3)
    # ---- roofline of the dominant kernel, the low-res scorer (rank 0's block).  What
    # binds it is one fp64 add per (candidate, point) in point order;
    # the byte views SURVEY 8(d) asks for (8 B per (candidate, point) against HBM and
    # against shared memory) are kept as secondary keys.
    pts = (offs[1:] - offs[:-1])[pa].astype(np.int64)
    adds_all = float((st[:, 0] * pts).sum())                  # whole batch, one step
    if world > 1:  # rank 0 timed its own block only
        w = m.pair_weights(pa, leg.rmin, leg.rmax, offs)
        order, bounds = D.partition_pairs(pb, w, world)
        mine = order[bounds[0]:bounds[1]]
        adds = float((st[mine, 0] * pts[mine]).sum()) * args.steps
    else:
        adds = adds_all * args.steps
    low_s = acc["low"] * 1e-3
    waves = max(1, m.last_timings()["waves"])
    peak, peak_kind = measured_peak_gbs()
    props = torch.cuda.get_device_properties(local)
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    n_sm = props.multi_processor_count
    fp64_peak = n_sm * 64 * sm_mhz * 1e6 / 1e9       # Gadd/s: 2 warp-DADD / clk / SM (scripts/micro/dadd_bench.cu)
    smem_peak = n_sm * 128 * sm_mhz * 1e6 / 1e9      # GB/s
    fp64_achieved = adds / low_s / 1e9
    traffic = ncu_traffic()
    tot_k = acc["tables"] + acc["low"] + acc["refine"] + acc["gather"]
    n_tables = args.scans if (world == 1 or main_strong) else args.scans  # per rank and step (weak: its own 2000)
    if main_strong:
        n_tables = args.scans / world
    dense_bytes = 8.0 * (m.w_high ** 2 + m.w_low ** 2) * n_tables * args.steps
    tab = {"ms_per_step": acc["tables"] / args.steps, "share": acc["tables"] / tot_k,
           "dense_layout_GBps": dense_bytes / (acc["tables"] * 1e-3) / 1e9,
           "dense_layout_note": "8 (W^2 + Wl^2) B per table, the reference's dense layout (SURVEY 8d); "
                                "the kernels only compute the cropped rectangle, so this is NOT a "
                                "bandwidth the hardware sees"}
    per_table = traffic.get("table_bytes_per_table") if traffic else None
    if per_table:
        actual = per_table * n_tables * args.steps / (acc["tables"] * 1e-3) / 1e9
        tab.update({"bound": "hbm", "actual_GBps": actual, "peak": peak, "frac": actual / peak,
                    "dram_bytes_per_table": per_table,
                    "min_bytes_per_table": traffic.get("table_min_bytes_per_table"),
                    "traffic_over_min": (per_table / traffic["table_min_bytes_per_table"]
                                         if traffic.get("table_min_bytes_per_table") else None),
                    "traffic_source": traffic.get("_source")})
    kernels = {
        "low_score_kernel": {"ms_per_step": acc["low"] / args.steps, "share": acc["low"] / tot_k},
        "refine_kernels": {"ms_per_step": acc["refine"] / args.steps, "share": acc["refine"] / tot_k,
                           "kernels": "refine_first + refine_spec + refine_finish",
                           "speculative_windows_per_step": m.last_timings()["spec_windows"]},
        "table_kernels": tab,
    }
    if world > 1:
        kernels["nccl_all_gather"] = {"ms_per_step": acc["gather"] / args.steps, "share": acc["gather"] / tot_k,
                                      "bytes_per_rank": int(48 * n_pairs / world)}
    cfg = workload_config(args.scans, world)
    if world > 1:
        cfg["parallelism"] = (
            f"one batch of {n_pairs} pairs over {len(offs) - 1} scans, partitioned by the library over "
            f"{world} ranks (sorted by target, blocks balanced by N_theta*N_t^2*N), ncclAllGather of the "
            f"48-byte records inside nautilus_csm_match_batch_sharded; every rank ends with all results")
        cfg["pairs_total"] = int(n_pairs)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "strong" if main_strong else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": cfg,
        "pairs_per_sec": n_pairs * args.steps / (ms * 1e-3),
        "e2e": {"value": e2e_value, "unit": UNIT,
                "h2d_bytes_per_step": leg.h2d, "d2h_bytes_per_step": int(48 * n_pairs),
                "ms_per_step": ms_e2e / args.steps,
                "ms_per_step_device_events": ms_e2e_dev / args.steps,
                "ms_per_step_host_wall": wall_e2e / args.steps,
                "path": ("nautilus_csm_upload_scans + nautilus_csm_match_batch" if world == 1 else
                         "per rank: nautilus_csm_partition + upload of the block's scans + "
                         "nautilus_csm_match_batch + nautilus_csm_gather")},
        "gpu_launches": acc["launches"],
        "clocks": clocks,
        "roofline": {"kernel": "low_score_kernel", "bound": "fp64_add", "achieved": fp64_achieved,
                     "peak": fp64_peak, "unit": "Gadd/s", "frac": fp64_achieved / fp64_peak,
                     "traffic": (traffic or {}).get("low_score_kernel_bytes_per_launch"),
                     "peak_source": f"{n_sm} SMs x 64 fp64 add/clk x {sm_mhz:.0f} MHz sampled in the timed "
                                    "region (pipe rate measured by scripts/micro/dadd_bench.cu; derived, "
                                    "not driver-measured)",
                     "launches": waves * args.steps,
                     "units_per_launch": adds / (waves * args.steps),
                     "ncu_fp64_pipe_pct": (traffic or {}).get("low_score_kernel_fp64_pipe_pct"),
                     "hbm_view": {"algorithmic_GBps": 8.0 * fp64_achieved, "peak": peak,
                                  "peak_source": peak_kind, "frac": 8.0 * fp64_achieved / peak},
                     "smem_view": {"algorithmic_GBps": 8.0 * fp64_achieved, "peak": smem_peak,
                                   "frac": 8.0 * fp64_achieved / smem_peak},
                     "note": "one fp64 add per (candidate, point), summed in point order like cc:66-77; "
                             "the staged table lives in shared memory and a patch is loaded once per run "
                             "of points that share it, so the SURVEY 8(d) byte model (8 B per (candidate, "
                             "point)) describes neither HBM (traffic ~ 0) nor shared memory: both views "
                             "exceed 1 and are informational only"},
        "kernels": kernels,
        "host_ms_per_step": {"plan_until_last_launch": acc["host_plan"] / args.steps,
                             "blocked_in_stream_sync": acc["host_wait"] / args.steps,
                             "collect_after_sync": acc["host_collect"] / args.steps},
        "step_wall_ms": {"resident": acc["step_wall_ms"], "e2e": acc_e2e["step_wall_ms"],
                         "e2e_upload": acc_e2e["upload_wall_ms"],
                         "resident_gpu_span": acc["gpu_span_ms"], "e2e_gpu_span": acc_e2e["gpu_span_ms"]},
    }
    if world > 1:
        line["gather_check"] = gather_check
        line["other_split"] = other
    if not args.no_cpu_baseline and world == 1:
        cores = host_cores()
        n_sample = min(n_pairs, 60 * cores)   # ~15 s of all-core CPU work
        idx = np.linspace(0, n_pairs - 1, n_sample).astype(np.int64)
        secs, kind, res = cpu_reference_run(xy, offs, pa, pb, idx, cores)
        cand = int(st[idx, 0].sum() + st[idx, 1].sum())
        same = bool(np.array_equal(res[1], x[idx]) and np.array_equal(res[2], y[idx]) and
                    np.array_equal(res[3], th[idx]))
        line["cpu_baseline"] = {
            "value": cand / secs, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{n_sample} of {n_pairs} pairs, one pair per thread, {secs:.1f} s",
            "pairs_per_sec": n_sample / secs, "poses_bit_identical_to_gpu": same}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
