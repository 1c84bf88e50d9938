"""nautilus_csm_table_normalize_noalloc: the allocation-free variant of
nautilus_csm_table_normalize (static __device__ scratch, no cudaMalloc), checked
through the C ABI against the existing entry point on the same tables.

The arithmetic is the same operations in the same order, so the bar is the one
tests/test_gpu_shim.py holds nautilus_csm_table_normalize to: bit-identical
(tolerance 0).  Two calls in a row share the static scratch: the second table is
SMALLER than the first, so partials left by the first call would change its result
if a call read slots it did not write.

Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from nautilus_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_CELLS = 67108864  # NAUTILUS_NOALLOC_MAX_CELLS, include/nautilus_csm.h


def reference_normalize(L, table):
    want = np.ascontiguousarray(table.copy())
    _lib.check(L.nautilus_csm_table_normalize(_lib.ptr(want), table.shape[1], table.shape[0], 0))
    return want


def noalloc_normalize(L, torch, table, in_place=False):
    src = torch.from_numpy(table).to("cuda:0")
    dst = src if in_place else torch.full_like(src, -7.0)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    stream = torch.cuda.current_stream().cuda_stream
    _lib.check(L.nautilus_csm_table_normalize_noalloc(
        C.c_void_p(src.data_ptr()), C.c_void_p(dst.data_ptr()), table.shape[1], table.shape[0], 0,
        C.c_void_p(stream)))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return dst.cpu().numpy()


def make_table(seed, h, w):
    rng = np.random.default_rng(seed)
    t = rng.random((h, w)) * 3.0 - 0.5
    t[rng.integers(0, h), rng.integers(0, w)] = 11.25   # one clear maximum
    t[rng.integers(0, h), rng.integers(0, w)] = -4.5    # and minimum, each in some block
    return t


def test_noalloc_normalize_matches_existing_entry_point_twice():
    import torch
    L = _lib.lib()
    # 1st call: 1500 x 1333 = 1 999 500 cells -> 489 scratch blocks, ragged last block
    # 2nd call: 61 x 97 = 5917 cells -> 2 blocks; slots 2.. still hold the first call's
    big, small = make_table(1, 1500, 1333), make_table(2, 61, 97)
    small *= 0.01  # its range lies strictly inside the first table's
    for table in (big, small):
        got = noalloc_normalize(L, torch, table)
        want = reference_normalize(L, table)
        assert got.min() == 0.0 and got.max() == 1.0
        assert np.array_equal(got, want), float(np.abs(got - want).max())
    # in place, constant table (m == M -> filled with 0) and an already-normalised one (untouched)
    for table in (np.full((33, 65), 2.5), np.linspace(0.0, 1.0, 4097 * 3).reshape(3, 4097)):
        got = noalloc_normalize(L, torch, table, in_place=True)
        assert np.array_equal(got, reference_normalize(L, table))


def test_noalloc_normalize_rejects_tables_beyond_the_static_scratch():
    import torch
    L = _lib.lib()
    x = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    p = C.c_void_p(x.data_ptr())
    rc = L.nautilus_csm_table_normalize_noalloc(p, p, 8193, 8192, 0, None)  # MAX_CELLS + 8192
    assert rc != 0
    msg = L.nautilus_csm_last_error().decode()
    assert str(MAX_CELLS) in msg and "maximum" in msg, msg
    assert L.nautilus_csm_table_normalize_noalloc(p, p, 0, 4, 0, None) != 0
    assert L.nautilus_csm_table_normalize_noalloc(None, p, 4, 4, 0, None) != 0
    # exactly at the limit the size check passes (not launched here: 512 MB per buffer)
    assert 8192 * 8192 == MAX_CELLS
