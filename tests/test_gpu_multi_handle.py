"""Several library handles in ONE process (the Julia single-process use the header advertises, include/dcegm.h:
"one handle per GPU"): kernel attributes such as the dynamic shared-memory opt-in are per device, so nothing
may be cached per process."""
import numpy as np
import pytest

from conftest import bm_param
from parity import assert_bitwise

pytestmark = pytest.mark.gpu


def _solve_all_models(s):
    from dcegm_jl_b200 import Param, _abi

    f = s.solve(_abi.MODEL_F, [bm_param()])                       # 107 KB of dynamic shared memory: needs the opt-in
    m = s.solve(_abi.MODEL_MINIMAL, [Param(par=dict(na=1500))])   # 72 KB: above the 48 KB default as well
    g = s.solve(_abi.MODEL_G, [Param(par=dict(na=100))])
    return f, m, g


def test_two_handles_one_device():
    from dcegm_jl_b200.solver import Solver

    with Solver(0) as a, Solver(0) as b:
        ra = _solve_all_models(a)
        rb = _solve_all_models(b)
        for x, y in zip(ra, rb):
            assert_bitwise(x, y)
            assert (x.status == 0).all()


def test_one_handle_per_device_in_one_process():
    import torch

    from dcegm_jl_b200.solver import Solver

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    with Solver(0) as a:
        ra = _solve_all_models(a)
        # the second handle is created after the first one has configured and launched its kernels on device 0
        with Solver(1) as b:
            rb = _solve_all_models(b)
            ra2 = _solve_all_models(a)  # and device 0 still works afterwards
    for x, y, z in zip(ra, rb, ra2):
        assert_bitwise(x, y)
        assert_bitwise(x, z)
        assert np.array_equal(x.status, y.status)
