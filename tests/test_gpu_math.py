"""The kernels' branch-free division / reciprocal (csrc/dc_math.h dc_div_fast, dc_rcp_fast) give the
IEEE-754 correctly rounded results on the operand ranges the solver uses — the CPU oracle's portable
mode uses the plain `/` there, so this is what makes the bitwise GPU-vs-oracle parity possible."""
import ctypes as C

import pytest

from dcegm_jl_b200 import _lib

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("emin,emax,seed", [(-40, 40, 1), (-90, 200, 2), (-3, 3, 3), (0, 1, 4)])
def test_fast_division_is_ieee(emin, emax, seed):
    lib = _lib.load()
    bad = (C.c_ulonglong * 2)()
    rc = lib.dcegm_debug_math_check(200_000_000, seed, emin, emax, bad)
    assert rc == 0
    assert bad[0] == 0 and bad[1] == 0, f"{bad[0]} divisions, {bad[1]} reciprocals differ from IEEE"
