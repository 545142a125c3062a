import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_c512():
    import numpy as np
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "cfg1_c512.npz")))


@pytest.fixture(scope="session")
def golden_c1024():
    import numpy as np
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "cfg2_c1024.npz")))
