"""Generates tests/golden/residual_golden.npz from the reference's own residual functors
(src/optimization/slam_residuals.h:64-177, compiled UNCHANGED into oracle/_ref and
evaluated through AutoDiffCostFunction::Evaluate; Ceres' Jet arithmetic and Eigen's
types come from oracle/shim, which restates their published algorithms because both
libraries are un-vendored).  Run in the build container:
    python tests/golden/make_residual_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import build, ref  # noqa: E402
from test_residuals import make_blocks  # noqa: E402

assert build.build_ref() is not None
out = {}
for kind in (0, 1):
    sp, tp, offs, p, q, ns, nt = make_blocks(40, seed=11 + kind, max_corr=30)
    res, js, jt = ref.residuals_eval(kind, sp, tp, offs, p, q, ns, nt)
    for name, v in (("sp", sp), ("tp", tp), ("offs", offs), ("p", p), ("q", q), ("ns", ns), ("nt", nt),
                    ("res", res), ("js", js), ("jt", jt)):
        out[f"k{kind}_{name}"] = v
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "residual_golden.npz"), **out)
print({k: v.shape for k, v in out.items()})
