import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# the FastAPI mirror builds VGG16(weights='imagenet') at import like the reference; there is no weight file on the
# test boxes, so the suite opts into the seeded random init explicitly (the default is to raise)
os.environ.setdefault("DCV_ALLOW_SYNTH", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


def rel_l2(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    d = np.linalg.norm((a - b).ravel())
    n = np.linalg.norm(b.ravel())
    return d / n if n > 0 else d


def unpack2(packed, n):
    p = np.asarray(packed, np.uint8)
    q = np.stack([p & 3, (p >> 2) & 3, (p >> 4) & 3, (p >> 6) & 3], axis=1).reshape(-1)
    return q[:n]


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    return load


@pytest.fixture(scope="session")
def engine():
    """Process-wide Engine on cuda:0.  Fails (does not skip) when the CUDA library cannot run."""
    from deconv_api_b200 import engine as E
    return E.default_engine(0)
