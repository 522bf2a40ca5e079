import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as ge
    p = ge.load_package()
    if not os.path.exists(p.LIB_PATH):
        p.build()
    return p


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def cos_and_maxabs(a, b):
    import numpy as np
    cos = (a * b).sum(1) / np.linalg.norm(a, axis=1) / np.linalg.norm(b, axis=1)
    return float(cos.min()), float(np.abs(a - b).max())
