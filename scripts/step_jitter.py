"""Step-time jitter study: resident-input steps with and without the NVML clock
sampler thread of bench.py running beside them."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from nautilus_b200 import CorrelativeScanMatcher

xy, offs, pa, pb = bench.make_batch(2000, seed=0)
m = CorrelativeScanMatcher(*bench.PARAMS, device=0)
st = torch.cuda.current_stream()
m.set_stream(st.cuda_stream)
m.upload_scans(xy, offs)
for _ in range(3):
    m.match_batch(pa, pb, *bench.ROT_WINDOW, with_stats=True)

def run(tag, n=8):
    ts = []
    for _ in range(n):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m.match_batch(pa, pb, *bench.ROT_WINDOW, with_stats=True)
        ts.append(1e3 * (time.perf_counter() - t0))
    t = m.last_timings()
    print(tag, " ".join(f"{v:.1f}" for v in ts), "| gpu span", round(t["ms_total"], 2), flush=True)

run("plain   ")
s = bench.ClockSampler(0)
s.start()
run("sampler ")
print(s.stop())
run("plain   ")
