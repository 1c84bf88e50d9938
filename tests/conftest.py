import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "csm_golden.npz"))


@pytest.fixture(scope="session")
def ref_lib():
    """The unmodified reference, when built here (/root/reference) or prebuilt (oracle/_ref)."""
    from oracle import build, ref
    if build.build_ref() is None:
        pytest.skip("oracle/_ref not available")
    return ref


def has_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
