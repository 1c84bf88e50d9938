"""BASELINE configs[0] for the GPU arm: single-pair GetTransformation through the C++
drop-in header, (a) one caller (latency), (b) T OpenMP threads sharing one matcher
object, which the library's combining queue turns into batches; the reference doing
the same (one pair per thread, its own code) is timed beside it.  One JSON line.

  python scripts/bench_concurrent.py [--threads 16] [--calls 64] [--cpu]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nautilus_b200 import _lib, synth  # noqa: E402


def build(exe):
    cmd = ["g++", "-std=c++17", "-O2", "-fopenmp", os.path.join(ROOT, "tests", "cpp", "concurrent_bench.cc"),
           "-I" + os.path.join(ROOT, "nautilus_b200", "cpp"), "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "oracle", "shim"), "-L" + _lib.LIB_DIR, "-lnautilus_csm_b200",
           "-Wl,-rpath," + _lib.LIB_DIR, "-o", exe]
    subprocess.run(cmd, check=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--threads", type=int, default=16)
    ap.add_argument("--calls", type=int, default=64)
    ap.add_argument("--pairs", type=int, default=64)
    ap.add_argument("--cpu", action="store_true")
    args = ap.parse_args()
    tmp = tempfile.mkdtemp()
    exe, blob = os.path.join(tmp, "concurrent_bench"), os.path.join(tmp, "pairs.bin")
    build(exe)
    pairs = [synth.make_pair(500 + k)[:2] for k in range(args.pairs)]
    with open(blob, "wb") as f:
        np.asarray([len(pairs)], np.int32).tofile(f)
        for a, b in pairs:
            np.asarray([len(a), len(b)], np.int32).tofile(f)
            a.tofile(f)
            b.tofile(f)
    r = subprocess.run([exe, blob, str(args.threads), str(args.calls)], capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        sys.exit(1)
    line = json.loads(r.stdout.strip().splitlines()[-1])
    line["config"] = ("single pair, CSM(30, 2, 0.3, 0.03), window +-0.5 rad, 1080-beam scans "
                      "(BASELINE configs[0]) through nautilus_b200/cpp/CorrelativeScanMatcher.h")
    if args.cpu:
        from oracle import ref
        cores = len(os.sched_getaffinity(0))
        k = min(len(pairs), 2 * cores)
        xy = np.concatenate([np.concatenate(p) for p in pairs[:k]])
        sizes = np.asarray([n for p in pairs[:k] for n in (len(p[0]), len(p[1]))])
        offs = np.zeros(2 * k + 1, np.int64)
        offs[1:] = np.cumsum(sizes)
        ids = np.arange(k, dtype=np.int32)
        secs, *_ = ref.RefMatcher(30, 2, 0.3, 0.03).match_batch(
            xy, offs, 2 * ids, 2 * ids + 1, np.full(k, -0.5), np.full(k, 0.5), threads=cores)
        line["cpu_baseline"] = {"value": k / secs, "unit": "pairs/s", "cores": cores, "kind": "reference",
                                "sample": f"{k} pairs, one pair per thread, {secs:.1f} s"}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
