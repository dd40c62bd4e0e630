import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    from oracle import rf_oracle

    z = np.load(os.path.join(GOLDEN_DIR, f"{name}.npz"))
    return rf_oracle.RawForest.load(z, "f_"), z["X"], z["P"], z["leaves"]


@pytest.fixture(params=GOLDEN_CASES)
def golden(request):
    raw, X, P, leaves = load_golden(request.param)
    return request.param, raw, X, P, leaves


def bits(a):
    """View float64 results as integers so comparisons are bit-for-bit (NaN-safe)."""
    return np.ascontiguousarray(a, dtype=np.float64).view(np.uint64)
