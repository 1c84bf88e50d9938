"""TEST INFRASTRUCTURE ONLY -- builds the two CPU checkers.

  oracle/_build/libcsm_oracle.so   plain-C restatement of the CSM path (csm_oracle.c)
  oracle/_build/libnn_oracle.so    restatement of the KD-tree / NN path (nn_oracle.cc)
  oracle/_ref/libnautilus_ref.so   the UNMODIFIED reference sources, compiled where
                                   they lie under /root/reference (only when that
                                   tree is present, i.e. in the build container;
                                   the GPU box uses the prebuilt file)

Flags for the reference follow its Release build (CMakeLists.txt:11,16:
-std=c++17 -O3 -DNDEBUG -fopenmp, no -march => no FMA contraction) plus
-Dcimg_display=0 because there is no X11.  The reference's own build system
(catkin/cmake + Eigen/glog/Ceres/ROS) cannot run here; its CSM and KD-tree
translation units compile directly against oracle/shim/.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = "/root/reference"
BUILD_DIR = os.path.join(HERE, "_build")
REF_DIR = os.path.join(HERE, "_ref")
PORT_LIB = os.path.join(BUILD_DIR, "libcsm_oracle.so")
NN_PORT_LIB = os.path.join(BUILD_DIR, "libnn_oracle.so")
REF_LIB = os.path.join(REF_DIR, "libnautilus_ref.so")


def _newer(target: str, sources: list[str]) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources if os.path.exists(s))


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError("oracle build failed")


def build_port(force: bool = False) -> str:
    srcs = [os.path.join(HERE, "csm_oracle.c")]
    nn_src = os.path.join(HERE, "nn_oracle.cc")
    os.makedirs(BUILD_DIR, exist_ok=True)
    if force or not _newer(PORT_LIB, srcs):
        _run(["gcc", "-std=c11", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall",
              *srcs, "-o", PORT_LIB, "-lm"])
    if force or not _newer(NN_PORT_LIB, [nn_src]):
        _run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall",
              nn_src, "-o", NN_PORT_LIB])
    return PORT_LIB


def build_ref(force: bool = False) -> str | None:
    """Returns the path, or None when /root/reference is absent and nothing is prebuilt."""
    csm_src = os.path.join(REF_ROOT, "third_party/csm/src/CorrelativeScanMatcher.cc")
    kd_src = os.path.join(REF_ROOT, "src/util/kdtree.cc")
    wrap = [os.path.join(HERE, f) for f in ("ref_wrap.cc", "ref_nn_wrap.cc", "ref_frontend_wrap.cc",
                                            "ref_nautilus_wrap.cc", "ref_residual_wrap.cc")]
    wrap = [s for s in wrap if os.path.exists(s)]
    if not os.path.exists(csm_src):
        return REF_LIB if os.path.exists(REF_LIB) else None
    shim = [os.path.join(dp, f) for dp, _, fs in os.walk(os.path.join(HERE, "shim")) for f in fs]
    if not force and _newer(REF_LIB, wrap + shim):
        return REF_LIB
    os.makedirs(REF_DIR, exist_ok=True)
    srcs = wrap + [csm_src]
    if os.path.exists(os.path.join(HERE, "ref_nn_wrap.cc")):
        srcs.append(kd_src)
    if os.path.exists(os.path.join(HERE, "ref_frontend_wrap.cc")):
        srcs.append(os.path.join(REF_ROOT, "third_party/csm/src/pointcloud_helpers.cc"))
    if os.path.exists(os.path.join(HERE, "ref_nautilus_wrap.cc")):
        srcs.append(os.path.join(REF_ROOT, "src/input/pointcloud_helpers.cc"))
        srcs.append(os.path.join(REF_ROOT, "src/loop_closure/lc_candidate_filter.cc"))
    _run(["g++", "-std=c++17", "-O3", "-DNDEBUG", "-fopenmp", "-Dcimg_display=0",
          "-fPIC", "-shared",
          "-I" + os.path.join(HERE, "shim"),
          "-I" + os.path.join(REF_ROOT, "third_party/csm/src"),
          "-I" + os.path.join(REF_ROOT, "src"),
          *srcs, "-o", REF_LIB])
    return REF_LIB


REF_GTESTS = os.path.join(REF_DIR, "ref_csm_gtests")


def build_ref_gtests(force: bool = False) -> str | None:
    """The reference's own unit-test file, third_party/csm/test/
    correlative_scan_matcher_tests.cc, compiled UNCHANGED where it lies against the
    product's drop-in header (nautilus_b200/cpp/CorrelativeScanMatcher.h) and linked with
    libnautilus_csm_b200.so.  The file is fed to g++ on stdin from tests/cpp/dropin/test/
    so that its `#include "../src/CorrelativeScanMatcher.h"` resolves to the drop-in
    instead of the reference header beside it.  Returns the binary's path, or None when
    the reference tree is absent and nothing is prebuilt."""
    repo = os.path.dirname(HERE)
    src = os.path.join(REF_ROOT, "third_party/csm/test/correlative_scan_matcher_tests.cc")
    lib_dir = os.path.join(repo, "nautilus_b200", "lib")
    if not os.path.exists(src):
        return REF_GTESTS if os.path.exists(REF_GTESTS) else None
    deps = [os.path.join(repo, "nautilus_b200", "cpp", "CorrelativeScanMatcher.h"),
            os.path.join(repo, "include", "nautilus_csm.h"),
            os.path.join(repo, "tests", "cpp", "gtest_shim", "gtest", "gtest.h")]
    if not force and _newer(REF_GTESTS, deps):
        return REF_GTESTS
    os.makedirs(REF_DIR, exist_ok=True)
    shim = os.path.join(repo, "tests", "cpp", "gtest_shim")
    obj = os.path.join(REF_DIR, "ref_csm_gtests.o")
    cmd = ["g++", "-std=c++17", "-O1", "-DNAUTILUS_CSM_NO_CIMG", "-x", "c++", "-c", "-", "-o", obj,
           "-I" + shim, "-I" + os.path.join(HERE, "shim"), "-I" + os.path.join(repo, "include")]
    with open(src, "rb") as f:
        r = subprocess.run(cmd, stdin=f, capture_output=True, text=False,
                           cwd=os.path.join(repo, "tests", "cpp", "dropin", "test"))
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stderr.decode())
        raise RuntimeError("reference gtest build failed")
    _run(["g++", "-std=c++17", "-O1", obj, os.path.join(shim, "gtest_main.cc"), "-I" + shim,
          "-L" + lib_dir, "-lnautilus_csm_b200", "-Wl,-rpath," + lib_dir, "-o", REF_GTESTS])
    os.remove(obj)
    return REF_GTESTS


def build_all(force: bool = False) -> None:
    build_port(force)
    build_ref(force)
    if os.path.exists(os.path.join(os.path.dirname(HERE), "nautilus_b200", "lib",
                                   "libnautilus_csm_b200.so")):
        build_ref_gtests(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv)
    print(PORT_LIB, os.path.exists(PORT_LIB))
    print(REF_LIB, os.path.exists(REF_LIB))
