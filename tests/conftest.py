import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def literals():
    """Known-answer matrices transcribed from the reference's tests by tests/golden/make_golden.py."""
    with open(os.path.join(ROOT, "tests", "golden", "reference_literals.json")) as f:
        raw = json.load(f)
    return {k: v["scale"] * np.asarray(v["rows"], dtype=np.float64) for k, v in raw.items()}
