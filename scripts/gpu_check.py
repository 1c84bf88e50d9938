"""First-contact GPU check: tables and matches against the oracle, with timings.
Run on the GPU box:  python scripts/gpu_check.py [n_pairs]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from nautilus_b200 import CorrelativeScanMatcher, synth  # noqa: E402
from oracle import port  # noqa: E402


def rel_err(got, want):
    return np.abs(got - want) / np.maximum(np.abs(want), 1e-10)


def main():
    n_pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    pm = port.PortMatcher(30, 2, 0.3, 0.03)
    a0, b0, _ = synth.make_pair(0)
    ph = pm.highres(b0)
    pl = pm.lowres(ph)
    for margin in (-1, 0):
        m = CorrelativeScanMatcher(30, 2, 0.3, 0.03, margin_cells=margin)
        t0 = time.time()
        hi, lo = m.GetLookupTables(b0)
        dt = time.time() - t0
        eh, el = rel_err(hi.values, ph.values), rel_err(lo.values, pl.values)
        print(f"margin={margin}: hi bit-exact={np.array_equal(hi.values, ph.values)} "
              f"max rel err hi={eh.max():.3e} low={el.max():.3e} "
              f"low bit-exact={np.array_equal(lo.values, pl.values)} ({dt*1e3:.1f} ms)")
        bad = 0
        for seed in range(n_pairs):
            a, b, _ = synth.make_pair(seed)
            got, st = m.GetTransformation(a, b, -0.5, 0.5, with_stats=True)
            want, wst = pm.match(a, b, -0.5, 0.5, with_stats=True)
            ok = (got[1][0][0], got[1][0][1], got[1][1]) == (want[1], want[2], want[3])
            ok_s = abs(got[0] - want[0]) <= 1e-5 * abs(want[0])
            ok_st = tuple(st) == tuple(int(v) for v in wst)
            if not (ok and ok_s and ok_st):
                bad += 1
                print("  MISMATCH seed", seed, got, st, want, wst)
        print(f"margin={margin}: {n_pairs - bad}/{n_pairs} pairs match; timings", m.last_timings())
        m.close()

    # batch
    xy, offs = synth.make_scan_set(100, group=5)
    pa, pb = synth.make_pair_list(100, group=5)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m.upload_scans(xy, offs)
    for it in range(3):
        t0 = time.time()
        score, x, y, th, st = m.match_batch(pa, pb, -0.5, 0.5, with_stats=True)
        dt = time.time() - t0
        print(f"batch {len(pa)} pairs: {dt*1e3:.1f} ms wall, timings {m.last_timings()}, "
              f"pops mean {st[:,2].mean():.2f}")
    # check a few batch results against the oracle
    bad = 0
    for k in range(0, len(pa), 37):
        ca = xy[offs[pa[k]]:offs[pa[k] + 1]]
        cb = xy[offs[pb[k]]:offs[pb[k] + 1]]
        want, wst = pm.match(ca, cb, -0.5, 0.5, with_stats=True)
        if (x[k], y[k], th[k]) != (want[1], want[2], want[3]) or \
                tuple(st[k]) != tuple(int(v) for v in wst):
            bad += 1
            print("  BATCH MISMATCH", k, (score[k], x[k], y[k], th[k]), st[k], want, wst)
    print("batch spot-check mismatches:", bad)
    t0 = time.time()
    cs = m.build_tables(np.arange(100, dtype=np.int32))
    print(f"build_tables 100: {(time.time()-t0)*1e3:.1f} ms, checksum {cs:.6f}, {m.last_timings()}")


if __name__ == "__main__":
    main()
