"""Table-stage time of the bench batch against the crop margin (cells)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from nautilus_b200 import CorrelativeScanMatcher

xy, offs, pa, pb = bench.make_batch(2000, seed=0)
for margin in (128, 112, 96, 80):
    m = CorrelativeScanMatcher(*bench.PARAMS, device=0, margin_cells=margin)
    m.upload_scans(xy, offs)
    for _ in range(4):
        out = m.match_batch(pa, pb, *bench.ROT_WINDOW, with_stats=True)
    t = m.last_timings()
    print(margin, {k: round(t[k], 2) for k in ("ms_total", "ms_tables", "ms_low", "ms_refine")}, flush=True)
    m.close()
