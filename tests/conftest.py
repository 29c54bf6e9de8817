import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o

    o.lib()
    return o


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    return np.load(os.path.join(ROOT, "tests", "golden", "matlab_retirement_na500.npz"))


def ftest_param(**over):
    """parameters of test/F_test.jl:34-40"""
    from dcegm_jl_b200 import Param

    d = dict(na=500, beta=1 / 1.05, sigma=0.35, R=1.05, delta=0.0)
    d["lambda"] = 0.000002
    d.update(over)
    return Param(par=d)


def bm_param(**over):
    """parameters of src/bench.jl:18-22"""
    from dcegm_jl_b200 import Param

    d = dict(na=500, beta=0.95, sigma=0.35, R=1.05)
    d["lambda"] = 0.000002
    d.update(over)
    return Param(par=d)
