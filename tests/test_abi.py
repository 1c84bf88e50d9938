"""CPU tests of the C-ABI boundary: the library loads, exports every symbol
include/nautilus_csm.h declares, fails loudly without a GPU, and its host-side
index maps agree with the oracle."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, has_gpu
from nautilus_b200 import _lib
from oracle import port


@pytest.fixture(scope="module")
def L():
    _lib.build()
    return _lib.lib()


def test_exports_every_declared_symbol(L):
    header = open(os.path.join(ROOT, "include", "nautilus_csm.h")).read()
    declared = set(re.findall(r"\b(nautilus_(?:csm|nn|lc|residuals|host)_\w+)\s*\(", header))
    assert declared, "no declarations found"
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name


def test_version(L):
    assert b"sm_100a" in L.nautilus_csm_version()


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU failure path")
def test_create_fails_loudly_without_gpu(L):
    p = _lib.Params(30, 2, 0.3, 0.03, 0, 0, 0, 0, 0, 0)
    h = C.c_void_p()
    rc = L.nautilus_csm_create(C.byref(p), C.byref(h))
    assert rc == -2 and not h.value
    assert b"CUDA" in L.nautilus_csm_last_error() or b"device" in L.nautilus_csm_last_error()
    from nautilus_b200 import CorrelativeScanMatcher, CsmError
    with pytest.raises(CsmError):
        CorrelativeScanMatcher(30, 2, 0.3, 0.03)


def test_create_rejects_bad_params(L):
    h = C.c_void_p()
    for bad in [(0, 2, 0.3, 0.03), (30, 2, 0, 0.03), (30, 2, 0.3, -1), (30, 0, 0.3, 0.03)]:
        p = _lib.Params(*bad, 0, 0, 0, 0, 0, 0)
        assert L.nautilus_csm_create(C.byref(p), C.byref(h)) == -1
    assert L.nautilus_csm_create(None, C.byref(h)) == -1


def test_cell_map_matches_oracle(L):
    rng = np.random.default_rng(0)
    O = port.lib()
    for res, dim in [(0.03, 2000), (0.3, 200), (1.0, 8), (0.01, 6000)]:
        vals = np.concatenate([rng.uniform(-35, 35, 2000), np.arange(-40, 40) * res,
                               [-30.0, 30.0, 29.999999, -30.000001, 0.0, -0.0]]).astype(np.float32)
        for v in vals:
            assert L.nautilus_csm_cell(res, dim, v) == O.oracle_cell(v, res, dim), (res, v)


def test_rotation_window_matches_oracle(L):
    pm = port.PortMatcher(30, 2, 0.3, 0.03)
    rng = np.random.default_rng(1)
    for _ in range(200):
        ra, rb = rng.uniform(-7, 7, 2)
        rr = rng.uniform(0.01, 6.2)
        lo, hi = C.c_double(), C.c_double()
        L.nautilus_csm_rotation_window(ra, rb, rr, C.byref(lo), C.byref(hi))
        assert (lo.value, hi.value) == pm.rotation_window(ra, rb, rr)


def test_cell_fast_equals_division_form(L):
    """The kernels find cells through float thresholds instead of a double
    division; the host twin of that map must agree with convertX everywhere:
    random values, every cell boundary and its float neighbours, far outside."""
    rng = np.random.default_rng(2)
    for res, dim in [(0.03, 2000), (0.3, 200), (0.01, 6000), (0.05, 160), (0.5, 16), (1.0, 8)]:
        half_extent = res * dim / 2
        edges = (np.arange(-dim // 2 - 2, dim // 2 + 3) * res).astype(np.float32)
        near = np.concatenate([edges] + [np.nextafter(edges, np.float32(s * np.inf)) for s in (-1, 1)])
        for _ in range(2):
            near = np.concatenate([near, np.nextafter(near, np.float32(-np.inf)),
                                   np.nextafter(near, np.float32(np.inf))])
        vals = np.concatenate([
            near, rng.uniform(-1.2 * half_extent, 1.2 * half_extent, 20000).astype(np.float32),
            np.asarray([0.0, -0.0, 1e-30, -1e-30, 1e9, -1e9, 3e38, -3e38], np.float32)])
        for v in np.unique(vals):
            want = L.nautilus_csm_cell(res, dim, v)
            want = want if 0 <= want < dim else -1
            assert L.nautilus_csm_cell_fast(res, dim, v) == want, (res, dim, float(v))
