"""Effect of the table crop margin (NautilusCsmParams::margin_cells) on table values
and table-build time, against the uncropped grid (margin -1, bit-identical to the
reference).  Run on a B200:  python scripts/margin_study.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nautilus_b200 import CorrelativeScanMatcher, synth

n = 1000
xy, offs = synth.make_scan_set(n, group=5, seed=0)
ids = np.arange(n, dtype=np.int32)
full = CorrelativeScanMatcher(30, 2, 0.3, 0.03, margin_cells=-1)
ref = [full.GetLookupTables(xy[offs[i]:offs[i + 1]]) for i in range(4)]
for margin in (72, 96, 112, 128, 144, 160):
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03, margin_cells=margin)
    worst, ndiff, nabove = 0.0, 0, 0
    for i in range(4):
        hi, lo = m.GetLookupTables(xy[offs[i]:offs[i + 1]])
        a, b = ref[i][0].values, hi.values
        above = np.maximum(a, b) > 1e-10           # cells the lookup does not floor
        d = np.abs(a - b)[above] / a[above]
        worst = max(worst, float(d.max()))
        ndiff += int((a[above] != b[above]).sum()); nabove += int(above.sum())
    m.upload_scans(xy, offs)
    m.build_tables(ids[:64])
    t = []
    for _ in range(3):
        m.build_tables(ids)
        t.append(m.last_timings()["ms_tables"])
    print(f"margin {margin:4d}: max rel diff above floor {worst:.3e}, cells differing {ndiff}/{nabove}, "
          f"{np.median(t) / n * 1e3:.2f} us/table", flush=True)
    m.close()
