"""One batch partitioned over two GPUs THROUGH THE C ABI (no torch): two processes,
one per GPU, each creates a handle with its world_rank / world_size, the NCCL unique
id travels through a file, and nautilus_csm_match_batch_sharded must return on every
rank exactly what one GPU returns for the whole batch (poses, scores and counters are
deterministic).  Also the explicit form: nautilus_csm_partition -> local
nautilus_csm_match_batch -> nautilus_csm_gather.  Needs two GPUs (gpurun --gpus 2)."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

WORKER = r'''
import os, sys, time
import numpy as np
sys.path.insert(0, os.environ["REPO_ROOT"])
from nautilus_b200 import CorrelativeScanMatcher, distributed as D, synth

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
id_path = os.environ["ID_PATH"]
xy, offs = synth.make_scan_set(60, group=5, seed=17, n_beams=360)
pa, pb = synth.make_pair_list(60, group=5)
perm = np.random.default_rng(2).permutation(len(pa))
pa, pb = pa[perm], pb[perm]
rmin = np.where(np.arange(len(pa)) % 3 == 0, -0.3, -0.5)
rmax = np.where(np.arange(len(pa)) % 3 == 0, 0.4, 0.5)

m = CorrelativeScanMatcher(30, 2, 0.3, 0.03, device=rank, world_rank=rank, world_size=world)
if rank == 0:
    uid = m.comm_unique_id()
    with open(id_path + ".tmp", "wb") as f:
        f.write(uid)
    os.replace(id_path + ".tmp", id_path)
else:
    for _ in range(600):
        if os.path.exists(id_path):
            break
        time.sleep(0.1)
    uid = open(id_path, "rb").read()
m.comm_init(uid)
m.upload_scans(xy, offs)
# first, on the fresh handle: blocks of very different size through the explicit form.  The
# smaller block's result buffer has to grow to the all-gather's send size (the largest
# block) without losing its records.
# (the pair list repeated 200 times: record buffers of 0.1 - 0.5 MB, which the allocator does
# not hand back in place the way it does a few KB)
REP = 200
pa_t, pb_t, rmin_t, rmax_t = (np.tile(v, REP) for v in (pa, pb, rmin, rmax))
order_u = np.argsort(pb_t, kind="stable").astype(np.int32)
bounds_u = np.asarray([0, len(pa_t) // 4, len(pa_t)], np.int64)
mine_u = order_u[bounds_u[rank]:bounds_u[rank + 1]]
m.match_batch(pa_t[mine_u], pb_t[mine_u], rmin_t[mine_u], rmax_t[mine_u], with_stats=True)
got_u = m.gather(order_u, bounds_u, with_stats=True)
got = m.match_batch_sharded(pa, pb, rmin, rmax, with_stats=True)
t = m.last_timings()
# the explicit form
w = m.pair_weights(pa, rmin, rmax, offs)
order, bounds = D.partition_pairs(pb, w, world)
mine = order[bounds[rank]:bounds[rank + 1]]
local = m.match_batch(pa[mine], pb[mine], rmin[mine], rmax[mine], with_stats=True)
got2 = m.gather(order, bounds, with_stats=True)
m.close()
# one GPU, whole batch
m1 = CorrelativeScanMatcher(30, 2, 0.3, 0.03, device=rank)
m1.upload_scans(xy, offs)
want = m1.match_batch(pa, pb, rmin, rmax, with_stats=True)
m1.close()
for u, v, z, zu in zip(got, want, got2, got_u):
    assert np.array_equal(u, v) and np.array_equal(z, v)
    assert np.array_equal(zu, np.tile(v, (REP,) + (1,) * (v.ndim - 1))), "unequal blocks: gathered records differ"
for u, v in zip(local, want):
    assert np.array_equal(u, v[mine])
assert 0 < len(mine) < len(pa)
print("rank", rank, "ok", len(mine), "of", len(pa), "ms_gather", t["ms_gather"])
'''


def test_sharded_match_two_gpus(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, REPO_ROOT=ROOT, WORLD_SIZE="2", ID_PATH=str(tmp_path / "nccl_id"))
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(2)]
    outs = [p.communicate(timeout=600)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, o
        assert f"rank {r} ok" in o


def test_sharded_world_size_one_is_match_batch():
    """With world_size 1 the sharded call is nautilus_csm_match_batch (no NCCL bound)."""
    import numpy as np
    from nautilus_b200 import CorrelativeScanMatcher, synth
    xy, offs = synth.make_scan_set(20, group=5, seed=3, n_beams=240)
    pa, pb = synth.make_pair_list(20, group=5)
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    m.upload_scans(xy, offs)
    a = m.match_batch(pa[::-1].copy(), pb[::-1].copy(), -0.5, 0.5, with_stats=True)
    b = m.match_batch_sharded(pa[::-1].copy(), pb[::-1].copy(), -0.5, 0.5, with_stats=True)
    for u, v in zip(a, b):
        assert np.array_equal(u, v)
    m.close()
