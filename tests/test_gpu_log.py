"""The table kernels' arithmetic helpers against the operations they stand for
(nautilus_csm_math_check): the table-driven logarithm against CUDA's log() - which is
itself within 1 ulp of the correctly rounded value the reference's glibc log returns
(CorrelativeScanMatcher.cc:72) - and the reciprocal-based division of the normalisation
against the IEEE division (CImg.h:39515), bit for bit.  Tolerances written here:
log within 2 ulps of log(), 0 differing quotients."""
import ctypes as C

import pytest

from nautilus_b200 import _lib

pytestmark = pytest.mark.gpu


def test_log_and_division_helpers():
    L = _lib.lib()
    ulps, bad = C.c_int64(-1), C.c_int64(-1)
    _lib.check(L.nautilus_csm_math_check(20_000_000, 12345, 0, C.byref(ulps), C.byref(bad)))
    assert 0 <= ulps.value <= 2, ulps.value
    assert bad.value == 0, bad.value
