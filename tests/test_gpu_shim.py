"""The C++ drop-in header (nautilus_b200/cpp/CorrelativeScanMatcher.h) compiled
the way a reference caller would compile it (g++, Eigen on the include path --
here the oracle's Eigen stand-in, since Eigen itself is absent from the image)
and linked against libnautilus_csm_b200.so."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT
from nautilus_b200 import CorrelativeScanMatcher, _lib, synth


def build_shim_test(tmp_path):
    exe = str(tmp_path / "shim_test")
    cmd = ["g++", "-std=c++17", "-O1", "-fopenmp", os.path.join(ROOT, "tests", "cpp", "shim_test.cc"),
           "-I" + os.path.join(ROOT, "nautilus_b200", "cpp"), "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "oracle", "shim"), "-L" + _lib.LIB_DIR,
           "-lnautilus_csm_b200", "-Wl,-rpath," + _lib.LIB_DIR, "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_shim_compiles_and_links(tmp_path):
    _lib.build()
    build_shim_test(tmp_path)


@pytest.mark.gpu
def test_shim_runs_reference_unit_tests_and_matches(tmp_path):
    exe = build_shim_test(tmp_path)
    a, b, _ = synth.make_pair(5, n_beams=360)
    blob = tmp_path / "pair.bin"
    with open(blob, "wb") as f:
        np.asarray([len(a), len(b)], np.int32).tofile(f)
        a.tofile(f)
        b.tofile(f)
    r = subprocess.run([exe, str(blob)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "SHIM OK" in r.stdout
    m = CorrelativeScanMatcher(30, 2, 0.3, 0.03)
    want = m.GetTransformation(a, b, -0.5, 0.5)
    want5 = m.GetTransformationRestricted(a, b, 0.1, 0.3, np.pi / 2)
    got = {ln.split()[0]: [float(v) for v in ln.split()[1:]] for ln in r.stdout.splitlines()
           if ln.startswith(("MATCH", "TABLEMAX", "UNC2", "STATS", "BYHAND", "OMPFIRST"))}
    for key, w in (("MATCH", want), ("MATCH5", want5)):
        assert got[key][0] == w[0]
        assert np.float32(got[key][1]) == w[1][0][0] and np.float32(got[key][2]) == w[1][0][1]
        assert np.float32(got[key][3]) == w[1][1]
    assert got["TABLEMAX"] == [1.0, 2000.0]
    # LookupTable::GaussianBlur / normalize on a hand-made table (dense recursion) against
    # the batch kernels' cropped table: <= 1e-5 relative, maximum exactly 1
    assert got["BYHAND"][1] <= 1e-5 and got["BYHAND"][2] == 1.0
    assert got["OMPFIRST"] == [1.0]  # concurrent first use of one matcher object
    u2 = m.GetUncertaintyMatrix(a, b, 0.125)
    assert np.array_equal(np.asarray(got["UNC2"], np.float32), u2.ravel())
    cond, scale = m.GetLocalUncertaintyStats([b], a)
    assert abs(got["STATS"][0] - cond) <= 1e-4 * abs(cond) and abs(got["STATS"][1] - scale) <= 1e-4 * abs(scale)
    m.close()


# ---- the reference's own gtest file, compiled UNCHANGED against the drop-in header ----
def test_reference_gtests_compile_against_dropin():
    """third_party/csm/test/correlative_scan_matcher_tests.cc (10 TESTs) builds against
    nautilus_b200/cpp/CorrelativeScanMatcher.h with the gtest stand-in; needs the
    reference tree (build container)."""
    from oracle import build
    _lib.build()
    exe = build.build_ref_gtests()
    if exe is None:
        pytest.skip("/root/reference not available and no prebuilt binary")
    assert os.path.exists(exe)


@pytest.mark.gpu
def test_reference_gtests_run_on_gpu():
    """Runs the binary built above (prebuilt into oracle/_ref/ in the build container,
    it travels to the GPU box): all 10 reference TESTs must pass on the GPU library."""
    from oracle import build
    exe = build.build_ref_gtests()
    if exe is None:
        pytest.skip("oracle/_ref/ref_csm_gtests not built")
    r = subprocess.run([exe], capture_output=True, text=True,
                       env=dict(os.environ, LD_LIBRARY_PATH=_lib.LIB_DIR + ":" + os.environ.get("LD_LIBRARY_PATH", "")))
    assert r.returncode == 0, r.stdout + r.stderr
    assert "10 tests ran, 0 failed" in r.stdout, r.stdout
