import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_clips():
    out = {}
    for name in ("29.1", "58.2", "116.4"):
        with np.load(os.path.join(GOLDEN, f"clip_{name}.npz")) as z:
            out[name] = {k: z[k] for k in z.files}
    return out


@pytest.fixture(scope="session")
def me():
    import motion_estimation_b200
    return motion_estimation_b200
