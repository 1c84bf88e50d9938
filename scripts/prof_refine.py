import sys, os
import numpy as np
sys.path.insert(0, "/root/repo")
from nautilus_b200 import CorrelativeScanMatcher, synth
n = 2000
xy, offs = synth.make_scan_set(n, group=5, seed=0)
pa, pb = synth.make_pair_list(n, group=5)
m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
m.upload_scans(xy, offs)
m.match_batch(pa[:100], pb[:100], -0.5, 0.5)
s, x, y, th, st = m.match_batch(pa, pb, -0.5, 0.5, with_stats=True)
pops = st[:, 2].astype(float); tot = s; words = st[:, 0].astype(float); gather = st[:, 1].astype(float)
print("total cycles/pair mean", tot.mean(), "max", tot.max(), "sum", tot.sum())
print("per pop: total", tot.sum() / pops.sum(), "words", words.sum() / pops.sum(), "gather", gather.sum() / pops.sum())
i = int(np.argmax(tot)); print("slowest pair: pops", pops[i], "cycles", tot[i], "per pop", tot[i] / pops[i], "words", words[i] / pops[i], "gather", gather[i] / pops[i])
one = pops == 1
print("1-pop pairs: total", tot[one].mean(), "words", words[one].mean(), "gather", gather[one].mean())
print(m.last_timings())
