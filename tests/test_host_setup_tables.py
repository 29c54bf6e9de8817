"""The library's own host-side constructors (csrc/host_setup.cpp: quadrature nodes, normal quantile, Rouwenhorst chain,
grids) against independent implementations (numpy leggauss / hermgauss, scipy norm.ppf, exact rational arithmetic,
the binomial stationary law) — the same checks tests/test_oracle_basics.py runs on the oracle's twin of these
functions, so that a shared misreading of param.jl / utils.jl cannot hide behind GPU == oracle.  No GPU needed."""
import ctypes as C
import math
from fractions import Fraction

import numpy as np
import pytest
from scipy import stats

from dcegm_jl_b200 import _lib

_dp = C.POINTER(C.c_double)


def table(which, n, p0=0.0, p1=0.0, p2=0.0, iarg=0, n1=None):
    lib = _lib.load()
    a = np.zeros(n)
    b = np.zeros(n1 if n1 is not None else n)
    assert lib.dcegm_debug_host_table(which, n, p0, p1, p2, iarg, a.ctypes.data_as(_dp), b.ctypes.data_as(_dp)) == 0
    return a, b


def test_quadpoints_is_gauss_legendre():  # utils.jl:261-297
    for n in (5, 32):
        x, w = table(1, n)
        xr, wr = np.polynomial.legendre.leggauss(n)
        assert np.allclose(x, 0.5 + 0.5 * xr, rtol=0, atol=2e-15)
        assert np.allclose(w, 0.5 * wr, rtol=1e-13)


def test_normal_quantile():  # param.jl:195-196
    for p in [1e-12, 1e-6, 0.01, 0.046910077, 0.2307653, 0.5, 0.77, 0.953089923, 1 - 1e-9]:
        assert table(2, 1, p)[0][0] == pytest.approx(stats.norm.ppf(p), rel=2e-15, abs=1e-16)


def test_gausshermite():  # param.jl:279
    for n in (5, 7, 32):
        x, w = table(3, n)
        xr, wr = np.polynomial.hermite.hermgauss(n)
        assert np.allclose(x, xr, rtol=1e-14, atol=1e-15)
        assert np.allclose(w, wr, rtol=1e-13)


def test_rouwenhorst():  # utils.jl:178-197
    z, P = table(4, 5, 0.95, 0.0, 0.35, n1=25)
    P = P.reshape(5, 5)
    assert np.allclose(P.sum(1), 1.0, rtol=1e-14)
    nu = math.sqrt(4 / (1 - 0.95 ** 2)) * 0.35
    assert z[0] == pytest.approx(-nu) and z[-1] == pytest.approx(nu) and z[2] == pytest.approx(0.0, abs=1e-15)
    q = (0.95 + 1) / 2
    assert P[0, 0] == pytest.approx(q ** 4) and P[0, 4] == pytest.approx((1 - q) ** 4)
    pi = np.array([math.comb(4, k) for k in range(5)]) / 16  # stationary law: Binomial(n-1, 1/2)
    assert np.allclose(pi @ P, pi, atol=1e-14)


def test_scalegrid_and_linrange():  # utils.jl:200-238, param.jl:199
    g, _ = table(5, 100, 0.0, 50.0, iarg=2)
    assert g[0] == pytest.approx(0.0, abs=1e-14) and g[-1] == pytest.approx(50.0, rel=1e-14)
    t = np.linspace(math.log(math.log(1.0) + 1), math.log(math.log(51.0) + 1), 100)
    assert np.allclose(g, np.exp(np.exp(t) - 1) - 1, rtol=1e-13, atol=1e-14)
    r, _ = table(0, 500, 0.0, 50.0)
    for i in (1, 2, 249, 333, 498):
        assert r[i] == float(Fraction(50 * i, 499))  # Julia's range is correctly rounded
    assert r[0] == 0.0 and r[-1] == 50.0


def test_same_tables_as_the_oracle(oracle):
    """bit for bit: the GPU-vs-oracle parity tests rely on both sides receiving identical tables"""
    for n in (5, 32):
        assert np.array_equal(table(1, n)[0], oracle.quadpoints(n)[0]) and np.array_equal(table(1, n)[1], oracle.quadpoints(n)[1])
        assert np.array_equal(table(3, n)[0], oracle.gausshermite(n)[0])
    assert np.array_equal(table(0, 500, 0.0, 50.0)[0], oracle.linrange(0.0, 50.0, 500))
    assert np.array_equal(table(5, 100, 0.0, 50.0, iarg=2)[0], oracle.scalegrid(0.0, 50.0, 100, 2))
    z, P = table(4, 5, 0.3, 0.0, 0.2, n1=25)
    zo, Po = oracle.rouwenhorst(0.3, 0.0, 0.2, 5)
    assert np.array_equal(z, zo) and np.array_equal(P.reshape(5, 5), Po)
