"""Summarise ncu output into profiles/ (run here, no GPU needed).

  python scripts/ncu_summary.py <tag> <launches.csv> <full.ncu-rep> [bench.json]

writes profiles/<tag>_launches.csv (copy), profiles/<tag>_summary.md and updates
profiles/traffic.json (dram read+write bytes per launch, per kernel) which
bench.py reports as roofline.traffic."""
import csv
import json
import os
import shutil
import subprocess
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROF = os.path.join(ROOT, "profiles")

METRICS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "L1/shared data pipe %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "shared wavefronts"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "shared bank conflicts"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit rate %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active % of peak"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64 pipe %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__occupancy_limit_registers", "blocks/SM limit (registers)"),
    ("launch__occupancy_limit_shared_mem", "blocks/SM limit (shared mem)"),
    ("launch__grid_size", "grid size"),
]


def short(name):
    return name.split("(")[0].replace("csm::", "").replace("void ", "").strip()


def launches_table(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    d = defaultdict(list)
    for r in rows[1:]:
        d[short(r[ki])].append(float(r[vi].replace(",", "")) / 1e6)
    total = sum(sum(v) for v in d.values())
    out = ["| kernel | launches | total ms | share |", "|---|---|---|---|"]
    for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
        out.append(f"| {k} | {len(v)} | {sum(v):.2f} | {sum(v) / total:.3f} |")
    return "\n".join(out), {k: sum(v) / total for k, v in d.items()}


def full_table(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    seen, cols = {}, []
    for r in rows[2:]:
        k = short(r[ki])
        if k not in seen:
            seen[k] = r
            cols.append(k)
    out = ["| metric | " + " | ".join(cols) + " |", "|---|" + "---|" * len(cols)]
    traffic = {}
    for m, label in METRICS:
        if m not in hdr:
            continue
        i = hdr.index(m)
        vals = []
        for k in cols:
            v = seen[k][i]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:.4g}"
            except ValueError:
                pass
            vals.append(f"{v} {units[i]}".strip())
        out.append(f"| {label} | " + " | ".join(vals) + " |")
    ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    ifp = hdr.index("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active") \
        if "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active" in hdr else None
    for k in cols:
        r = seen[k]
        rd = float(r[ir].replace(",", "")) * scale.get(units[ir], 1)
        wr = float(r[iw].replace(",", "")) * scale.get(units[iw], 1)
        traffic[k] = {"read": rd, "write": wr, "total": rd + wr,
                      "fp64_pipe_pct": float(r[ifp].replace(",", "")) if ifp is not None else None}
    return "\n".join(out), traffic


def main():
    tag, launches, rep = sys.argv[1], sys.argv[2], sys.argv[3]
    bench = sys.argv[4] if len(sys.argv) > 4 else None
    n_tables = int(sys.argv[5]) if len(sys.argv) > 5 else 2000  # bench.py --scans N of the capture
    os.makedirs(PROF, exist_ok=True)
    shutil.copy(launches, os.path.join(PROF, f"{tag}_launches.csv"))
    lt, shares = launches_table(launches)
    ft, traffic = full_table(rep)
    # traffic.json is rewritten from THIS capture only (no stale kernels of older builds)
    tab = [k for k in traffic if any(t in k for t in ("raster_kernel", "xpass_kernel", "ypass_kernel",
                                                      "ypass2_kernel", "low_finalize_kernel",
                                                      "normalize_pool_kernel"))]
    # the kernel that writes every cell of every stored rectangle exactly once: the
    # normalise pass (one-run path) or run 2 of the two-run y-pass (template MODE 2)
    norm = [k for k in traffic if "normalize_pool_kernel" in k] or \
           [k for k in traffic if "ypass2_kernel" in k and k.rstrip().endswith(", 2>")]
    low = [k for k in traffic if "low_score_kernel" in k]
    out = {"_source": f"{tag}: ncu --set full --clock-control none, one launch per kernel of "
                      f"`python bench.py --scans {n_tables} --steps 1 --warmup 1 --no-cpu-baseline` "
                      f"({n_tables} tables, {5 * n_tables} pairs per launch)",
           "tables_per_launch": n_tables,
           "kernels_bytes_per_launch": {k: v["total"] for k, v in traffic.items()},
           "table_bytes_per_table": sum(traffic[k]["total"] for k in tab) / n_tables,
           # its write bytes are the write-once minimum of this layout
           "table_min_bytes_per_table": (traffic[norm[0]]["write"] / n_tables) if norm else None,
           "low_score_kernel_bytes_per_launch": traffic[low[0]]["total"] if low else None,
           "low_score_kernel_fp64_pipe_pct": traffic[low[0]]["fp64_pipe_pct"] if low else None}
    json.dump(out, open(os.path.join(PROF, "traffic.json"), "w"), indent=1)
    md = [f"# ncu summary {tag}", "",
          "Launch list: `ncu --metrics gpu__time_duration.sum --clock-control none` over "
          f"`python bench.py --scans {n_tables} --steps 1 --warmup 1 --no-cpu-baseline` "
          f"({5 * n_tables} pairs, {n_tables} tables; cold-cache, serialised: compare SHARES).", "", lt, "",
          "Full capture (`ncu --set full --clock-control none --import-source on`, one launch per "
          "kernel of the same command):", "", ft, ""]
    if bench and os.path.exists(bench):
        b = json.loads(open(bench).read().strip().splitlines()[-1])
        shutil.copy(bench, os.path.join(PROF, f"{tag}_bench.json"))
        ks = b.get("kernels", {})
        md += ["Full-size bench line of the same build (`python bench.py`, CUDA events, no profiler): "
               f"{b['value']:.4g} {b['unit']}, {b['ms_per_step']:.1f} ms/step.  Live shares: " +
               ", ".join(f"{k} {v['share']:.3f}" for k, v in ks.items()) + ".", ""]
    open(os.path.join(PROF, f"{tag}_summary.md"), "w").write("\n".join(md))
    print("\n".join(md))


if __name__ == "__main__":
    main()
