"""world_size-2 gloo test (CPU) of the multi-GPU host logic: partition by target,
per-rank matching, all_gather of the 48-byte records, scatter back to caller
order.  The per-rank matcher here is the CPU oracle (tests may use it); on GPUs
bench.py injects the CUDA matcher and an NCCL all_gather instead."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT
from nautilus_b200 import distributed as D


def _np_partition(b_ids, weights, world):
    """The partition rule restated in numpy: stable sort by target, cut at
    the first position whose running weight reaches r / world of the total."""
    b_ids = np.asarray(b_ids)
    n = len(b_ids)
    order = np.argsort(b_ids, kind="stable")
    cum = np.cumsum(np.asarray(weights, np.float64)[order]) if n else np.zeros(0)
    total = cum[-1] if n else 0.0
    bounds = [0] + [int(np.searchsorted(cum, total * r / world, side="left")) if n else 0
                    for r in range(1, world)] + [n]
    return order, np.maximum.accumulate(np.asarray(bounds, np.int64))


def test_c_partition_equals_numpy_rule():
    """nautilus_csm_partition is host code of the library: it runs here without a GPU."""
    rng = np.random.default_rng(0)
    for trial in range(100):
        n, world = int(rng.integers(0, 600)), int(rng.integers(1, 9))
        b = rng.integers(0, rng.integers(1, 50) if trial % 3 else 10 ** 6, size=n)  # dense / sparse ids
        w = rng.uniform(1, 5, size=n)
        o1, b1 = _np_partition(b, w, world)
        o2, b2 = D.partition_pairs(b, w, world)
        assert np.array_equal(o1, o2) and np.array_equal(b1, b2), trial


def test_c_pair_weights():
    from nautilus_b200 import _lib
    p = _lib.Params(30, 2, 0.3, 0.03, 0, 0, 0, 0, 0, 0, 0, 1)
    w = D.pair_weights_c(p, [0, 1], [-0.5, 0.0], [0.5, 2 * np.pi], [0, 1080, 2000])
    assert w.tolist() == [58 * 14 * 14 * 1080, 361 * 14 * 14 * 920]  # trip counts of cc:92-99


def test_partition_properties():
    rng = np.random.default_rng(0)
    b = rng.integers(0, 40, size=500)
    w = rng.uniform(1, 5, size=500)
    for world in (1, 2, 3, 8):
        order, bounds = D.partition_pairs(b, w, world)
        assert sorted(order.tolist()) == list(range(500))
        assert bounds[0] == 0 and bounds[-1] == 500 and (np.diff(bounds) >= 0).all()
        assert (np.diff(b[order]) >= 0).all()           # sorted by target
        loads = [w[order[bounds[r]:bounds[r + 1]]].sum() for r in range(world)]
        assert max(loads) <= w.sum() / world + w.max() + 1e-9
        # a target is split across at most two adjacent ranks
        owners = {}
        for r in range(world):
            for t in set(b[order[bounds[r]:bounds[r + 1]]].tolist()):
                owners.setdefault(t, []).append(r)
        assert all(len(v) <= 2 or world > len(v) for v in owners.values())
    order, bounds = D.partition_pairs([], [], 4)
    assert len(order) == 0 and bounds.tolist() == [0, 0, 0, 0, 0]


WORKER = r'''
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.environ["REPO_ROOT"])
from nautilus_b200 import distributed as D, synth
from oracle import port

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
params = (4.0, 1.0, 0.5, 0.05)
rng = np.random.default_rng(1)
clouds = [rng.uniform(-2, 2, size=(60, 2)).astype(np.float32) for _ in range(4)]
a_ids = np.asarray([0, 1, 2, 3, 1, 0, 3], np.int32)
b_ids = np.asarray([1, 0, 3, 2, 1, 2, 0], np.int32)
pm = port.PortMatcher(*params)

def match_fn(a, b, rmin, rmax):
    out = np.zeros(len(a), dtype=D.RECORD)
    for k in range(len(a)):
        r, st = pm.match(clouds[a[k]], clouds[b[k]], rmin[k], rmax[k], with_stats=True)
        out[k] = (r[0], r[1], r[2], r[3], 0, st[0], st[1], st[2])
    return out

def all_gather(local, counts):
    mx = int(counts.max())
    buf = torch.zeros(mx * 48, dtype=torch.uint8)
    raw = torch.from_numpy(np.frombuffer(local.tobytes(), dtype=np.uint8).copy())
    buf[:raw.numel()] = raw
    outs = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(outs, buf)
    parts = [np.frombuffer(o.numpy().tobytes(), dtype=D.RECORD)[:int(c)] for o, c in zip(outs, counts)]
    return np.concatenate(parts)

w = D.pair_weights(np.full(len(a_ids), 58), 4, [len(clouds[a]) for a in a_ids])
got = D.match_batch_sharded(match_fn, a_ids, b_ids, -0.5, 0.5, w, rank=rank, world_size=world,
                            all_gather=all_gather)
want = match_fn(a_ids, b_ids, np.full(7, -0.5), np.full(7, 0.5))
assert got.tobytes() == want.tobytes(), (rank, got, want)
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
'''


def test_sharded_match_gloo_world2(tmp_path):
    import subprocess
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, REPO_ROOT=ROOT, MASTER_ADDR="127.0.0.1", MASTER_PORT="29613",
               WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, o
        assert f"rank {r} ok" in o
