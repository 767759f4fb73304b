import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fun-asr-vllm_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def frontend_golden():
    import numpy as np

    return np.load(os.path.join(GOLDEN, "frontend_golden.npz"))


@pytest.fixture(scope="session")
def encoder_golden():
    import numpy as np

    return np.load(os.path.join(GOLDEN, "encoder_golden.npz"))
