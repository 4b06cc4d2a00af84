import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def shipped_image():
    """the reference's 2d-convolution.input (512x512, integer values 0..255) as float64"""
    import numpy as np
    z = np.load(os.path.join(ROOT, "tests", "golden", "conv_input_512.npz"))
    return z["image"].astype(np.float64)


@pytest.fixture(scope="session")
def hgb():
    """the product package with its library loaded; GPU tests go through the C ABI only"""
    import hg_c_benchmarks_b200 as H
    H.lib()
    return H
