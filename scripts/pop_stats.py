"""Distribution of best-first pops per pair on the bench batch (GPU)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nautilus_b200 import CorrelativeScanMatcher, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 400
xy, offs = synth.make_scan_set(n, group=5, seed=0)
pa, pb = synth.make_pair_list(n, group=5)
m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
m.upload_scans(xy, offs)
s, x, y, th, st = m.match_batch(pa, pb, -0.5, 0.5, with_stats=True)
pops = st[:, 2]
print("pairs", len(pops), "mean pops", pops.mean(), "max", pops.max())
for q in (50, 75, 90, 95, 99, 99.9):
    print("p%g" % q, np.percentile(pops, q))
h = np.bincount(np.minimum(pops, 40))
print("hist (capped at 40):", h.tolist())
print("n_high mean", st[:, 1].mean(), "total pops", pops.sum())
print(m.last_timings())
