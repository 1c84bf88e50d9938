"""bench.py -- candidate transforms scored per second for the loop-closure batch
(BASELINE.json configs[2]: 10k synthetic scan pairs, multi-res branch search).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch: build the lookup tables of
the batch's 2000 target scans and match its 10 000 (query, target) pairs.  With
N GPUs every rank owns its own batch (weak scaling, pairs are independent); the
only exchange is an NCCL all_gather of the 48-byte per-pair results.

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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--stage-cells", type=int, default=0, help="tuning: staged low-res table budget")
    return ap.parse_args()


def workload_config(n_scans, n_gpus):
    return {
        "workload": "loop-closure candidate batch (BASELINE configs[2])",
        "pairs_per_gpu": n_scans * GROUP, "scans_per_gpu": n_scans,
        "beams": 1080, "matcher": "CSM(range=30, trans=2, low=0.3, high=0.03)",
        "rotation_window_rad": list(ROT_WINDOW), "rot_step_deg": 1.0,
        "parallelism": f"pairs x{n_gpus} (independent batches, all_gather of results)",
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


def ncu_traffic_bytes(kernel):
    """dram read+write bytes per launch of `kernel` from the committed ncu capture
    (profiles/*_traffic.json, written by scripts/ncu_summary.py); None if absent."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(p)).get(kernel)
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
    sample = (f"{n_sample} of {args.scans * GROUP} pairs per step, one pair per thread, tables "
              f"rebuilt per pair as GetTransformation does")
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
def run_ours(args):
    import torch
    import torch.distributed as dist
    from nautilus_b200 import CorrelativeScanMatcher

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

    xy, offs, pa, pb = make_batch(args.scans, seed=rank)
    n_pairs = len(pa)
    m = CorrelativeScanMatcher(*PARAMS, device=local, stage_cells=args.stage_cells)
    stream = torch.cuda.current_stream()
    m.set_stream(stream.cuda_stream)

    # pinned host staging for the end-to-end leg
    xy_pin = torch.from_numpy(xy).pin_memory()
    xy_host = xy_pin.numpy()
    rec = torch.empty(n_pairs * m.RESULT_BYTES, dtype=torch.uint8, device=dev)
    gathered = torch.empty(world * rec.numel(), dtype=torch.uint8, device=dev) if world > 1 else None

    upload_ms = [0.0]

    def step(upload):
        if upload:
            tu = time.perf_counter()
            m.upload_scans(xy_host, offs)
            upload_ms[0] = round(1e3 * (time.perf_counter() - tu), 3)
        out = m.match_batch(pa, pb, *ROT_WINDOW, with_stats=True)
        if world > 1:  # gather only the per-pair (pose, score) records
            m.copy_results_dev(rec.data_ptr(), n_pairs)
            dist.all_gather_into_tensor(gathered, rec)
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(upload, steps):
        """K steps between CUDA events on the launching stream; max over ranks."""
        acc = {"tables": 0.0, "low": 0.0, "refine": 0.0, "launches": 0, "cand": 0, "low_cand_pts": 0,
               "step_wall_ms": [], "gpu_span_ms": [], "upload_wall_ms": []}
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(steps):
            ts = time.perf_counter()
            score, x, y, th, st = step(upload)
            acc["step_wall_ms"].append(round(1e3 * (time.perf_counter() - ts), 3))
            t = m.last_timings()
            acc["gpu_span_ms"].append(round(t["ms_total"], 3))
            if upload:
                acc["upload_wall_ms"].append(upload_ms[0])
            acc["tables"] += t["ms_tables"]; acc["low"] += t["ms_low"]; acc["refine"] += t["ms_refine"]
            acc["launches"] += t["launches"]
            acc["cand"] += int(st[:, 0].sum() + st[:, 1].sum())
        e1.record(stream)
        barrier()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        ms = e0.elapsed_time(e1)
        if world > 1:
            v = torch.tensor([ms, wall_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(v, op=dist.ReduceOp.MAX)
            ms, wall_ms = float(v[0]), float(v[1])
            c = torch.tensor([acc["cand"]], device=dev, dtype=torch.int64)
            dist.all_reduce(c)
            acc["cand_all"] = int(c[0])
        else:
            acc["cand_all"] = acc["cand"]
        return ms, wall_ms, acc, (score, x, y, th, st)

    m.upload_scans(xy_host, offs)
    # NVML start-up, the sampler thread's start and its first queries (which can
    # stall the driver for ~0.2 s) happen during the warm-up; only samples taken
    # inside the timed region are kept
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        step(False)
    sampler.samples = []
    ms, _, acc, last = timed(False, args.steps)          # inputs resident in HBM
    clocks = sampler.stop() if rank == 0 else None
    for _ in range(1):
        step(True)
    ms_e2e_dev, wall_e2e, acc_e2e, _ = timed(True, args.steps)  # host buffers, copies inside
    ms_e2e = max(ms_e2e_dev, wall_e2e)  # the call is synchronous: host time counts

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    score, x, y, th, st = last
    value = acc["cand_all"] / (ms * 1e-3)
    e2e_value = acc_e2e["cand_all"] / (ms_e2e * 1e-3)
    # roofline of the dominant kernel (low-res scorer): algorithmic bytes = one
    # 8-byte table read per (candidate, point)  [SURVEY 8(d): 8*N B per candidate]
    pts = (offs[1:] - offs[:-1])[pa].astype(np.int64)
    low_bytes = float((st[:, 0] * pts).sum() * 8) * args.steps
    waves = max(1, m.last_timings()["waves"])
    peak, peak_kind = measured_peak_gbs()
    achieved = low_bytes / (acc["low"] * 1e-3) / 1e9
    props = torch.cuda.get_device_properties(local)
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    smem_peak = props.multi_processor_count * 128 * sm_mhz * 1e6 / 1e9
    fp64_peak = props.multi_processor_count * 64 * sm_mhz * 1e6 / 1e9       # Gadd/s
    fp64_achieved = low_bytes / 8.0 / (acc["low"] * 1e-3) / 1e9             # one add per table read
    table_bytes = 8.0 * (m.w_high ** 2 + m.w_low ** 2) * args.scans * args.steps
    kernels = {
        "low_score_kernel": {"ms_per_step": acc["low"] / args.steps,
                             "share": acc["low"] / (acc["tables"] + acc["low"] + acc["refine"])},
        "refine_kernels": {"ms_per_step": acc["refine"] / args.steps,
                           "share": acc["refine"] / (acc["tables"] + acc["low"] + acc["refine"]),
                           "kernels": "refine_first + refine_spec + refine_finish",
                           "speculative_windows_per_step": m.last_timings()["spec_windows"]},
        "table_kernels": {"ms_per_step": acc["tables"] / args.steps,
                          "share": acc["tables"] / (acc["tables"] + acc["low"] + acc["refine"]),
                          "algorithmic_GBps": table_bytes / (acc["tables"] * 1e-3) / 1e9,
                          "frac_of_hbm_peak": table_bytes / (acc["tables"] * 1e-3) / 1e9 / peak},
    }
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.scans, world),
        "pairs_per_sec": world * n_pairs * args.steps / (ms * 1e-3),
        "e2e": {"value": e2e_value, "unit": UNIT,
                "h2d_bytes_per_step": int(xy.nbytes + offs.nbytes + pa.nbytes + pb.nbytes + 16 * n_pairs),
                "d2h_bytes_per_step": int(48 * n_pairs),
                "ms_per_step": ms_e2e / args.steps,
                "ms_per_step_device_events": ms_e2e_dev / args.steps,
                "ms_per_step_host_wall": wall_e2e / args.steps},
        "gpu_launches": acc["launches"],
        "clocks": clocks,
        "roofline": {"kernel": "low_score_kernel", "bound": "hbm", "achieved": achieved,
                     "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic_bytes("low_score_kernel"),
                     "peak_source": peak_kind, "launches": waves * args.steps,
                     "smem_peak": smem_peak, "smem_frac": achieved / smem_peak,
                     "fp64_add_peak": fp64_peak, "fp64_add_achieved": fp64_achieved,
                     "fp64_add_frac": fp64_achieved / fp64_peak,
                     "note": "algorithmic 8 B per (candidate, point), SURVEY 8(d).  The kernel "
                             "stages its table in shared memory and loads a patch once per run "
                             "of points that share it, so neither HBM (traffic ~0, frac > 1 "
                             "expected) nor shared memory (smem_peak = SMs x 128 B/clk x SM "
                             "clock) binds it: what remains is one fp64 add per (candidate, "
                             "point) in point order.  fp64_add_peak = SMs x 64 adds/clk x SM "
                             "clock (measured: scripts/micro/dadd_bench.cu, 2 warp-DADD/clk/SM), "
                             "in Gadd/s; fp64_add_frac is the binding fraction (DESIGN.md 3)"},
        "kernels": kernels,
        "step_wall_ms": {"resident": acc["step_wall_ms"], "e2e": acc_e2e["step_wall_ms"],
                         "e2e_upload": acc_e2e["upload_wall_ms"],
                         "resident_gpu_span": acc["gpu_span_ms"], "e2e_gpu_span": acc_e2e["gpu_span_ms"]},
    }
    if not args.no_cpu_baseline:
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
