"""Generates tests/golden/unc_golden.npz from the UNMODIFIED reference
(oracle/_ref): GetUncertaintyMatrix 3x3 / 2x2 (CorrelativeScanMatcher.cc:267-373).
Run in the build container only:  python tests/golden/make_unc_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from nautilus_b200 import synth  # noqa: E402
from oracle import build, ref  # noqa: E402


def main():
    assert build.build_ref() is not None, "/root/reference not available"
    out = {}
    std = (30.0, 2.0, 0.3, 0.03)
    rm = ref.RefMatcher(*std)
    u2, u3, rots = [], [], []
    for k in range(3):
        a, b, truth = synth.make_pair(40 + k, n_beams=360)
        out[f"a_{k}"], out[f"b_{k}"] = a, b
        rot = float(np.float32(truth[2]))
        rots.append(rot)
        u2.append(rm.uncertainty2(a, b, rot))
        if k < 2:
            u3.append(rm.uncertainty3(a, b))
    out["rot"] = np.asarray(rots)
    out["u2"] = np.asarray(u2)
    out["u3"] = np.asarray(u3)
    out["params"] = np.asarray(std)
    np.savez_compressed(os.path.join(HERE, "unc_golden.npz"), **out)
    print(out["u2"], out["u3"])


if __name__ == "__main__":
    main()
