"""the batched call sites as a real patch against the reference
(patches/solver_batched.patch: solver.cc / solver.h / default_config.lua / CMakeLists.txt
plus the new src/optimization/csm_batch.{h,cc}).  patches/check.sh applies it to
pristine copies of the reference files and type-checks the new translation unit
against the reference's real headers.  Needs /root/reference (build container)."""
import os
import subprocess

import pytest

from conftest import ROOT


def test_solver_patch_applies_and_new_unit_type_checks():
    if not os.path.exists("/root/reference/src/optimization/solver.cc"):
        pytest.skip("/root/reference not available")
    r = subprocess.run([os.path.join(ROOT, "patches", "check.sh")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "patch applies" in r.stdout
    assert "csm_batch.cc type-checks against the reference headers" in r.stdout


def test_solver_patch_touches_the_cited_call_sites():
    text = open(os.path.join(ROOT, "patches", "solver_batched.patch")).read()
    for needle in ("AddLCConstraints", "BuildOptimizationOverWindow", "ResolveWithConstraints(lc_matches)",
                   "csm_score_threshold", "csm_high_res=0.01", "nautilus_csm_match_batch",
                   "nautilus_nn_match_batch", "CONFIG_csm_batch_matching"):
        assert needle in text, needle
