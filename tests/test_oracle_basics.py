"""test/basics.jl and the host-side numerics of the oracle (quadrature, grids, utilities)."""
import math

import numpy as np
import pytest
from scipy import special, stats

from dcegm_jl_b200 import Param


def test_param_override():  # basics.jl:2-5
    p = Param(par={"delta": 0.9})
    assert p.delta == 0.9


def test_param_defaults_and_derived():  # param.json, param.jl:90-93
    p = Param()
    assert (p.na, p.ny, p.nT, p.nD, p.retage) == (100, 5, 25, 2, 26)
    assert p.gamma == 1.95 and p.lambda_ == 2e-6 and p.R == 1.0526315789473684
    assert p.neg_gamma == -1.95 and p.oneminusgamma == 1.0 - 1.95
    assert p.neg_oneover_gamma == -1.0 / 1.95


def test_param_no_type_conversion():  # param.jl:82-87: setfield! does not convert
    with pytest.raises(TypeError):
        Param(par={"na": 500.0})
    with pytest.raises(KeyError):
        Param(par={"nope": 1.0})


def test_ccp(oracle):  # basics.jl:14-20
    rng = np.random.default_rng(0)
    x = rng.random((2, 10))
    c = oracle.ccp(x[0], x[1], Param().lambda_)
    assert len(c) == 10 and c.min() >= 0.0 and c.max() <= 1.0


def test_logsum(oracle):  # basics.jl:21-28
    z = np.zeros(2)
    for lam in (Param().lambda_, 1.0):
        assert np.max(np.abs(oracle.logsum(z, z, lam) - lam * math.log(2))) < 1e-6
    # exactness properties used by the CUDA path: the winner's exp is exactly 1
    v1 = np.array([-3.0, 2.0]); v2 = np.array([-1.0, -40.0])
    assert np.array_equal(oracle.ccp(v1, v2, 2e-6), [0.0, 1.0])
    assert np.array_equal(oracle.logsum(v1, v2, 2e-6), [-1.0, 2.0])


def test_quadpoints_is_gauss_legendre(oracle):  # utils.jl:261-297
    for n in (5, 32):
        x, w = oracle.quadpoints(n)
        xr, wr = np.polynomial.legendre.leggauss(n)
        assert np.allclose(x, 0.5 + 0.5 * xr, rtol=0, atol=2e-15)
        assert np.allclose(w, 0.5 * wr, rtol=1e-13)


def test_norminv(oracle):  # quantile(Normal(0,1), .) param.jl:195-196
    for p in [1e-12, 1e-6, 0.01, 0.046910077, 0.2307653, 0.5, 0.77, 0.953089923, 1 - 1e-9]:
        assert oracle.norminv(p) == pytest.approx(stats.norm.ppf(p), rel=2e-15, abs=1e-16)


def test_gausshermite(oracle):  # FastGaussQuadrature.gausshermite
    for n in (5, 7, 32):
        x, w = oracle.gausshermite(n)
        xr, wr = np.polynomial.hermite.hermgauss(n)
        assert np.allclose(x, xr, rtol=1e-14, atol=1e-15)
        assert np.allclose(w, wr, rtol=1e-13)


def test_rouwenhorst(oracle):  # utils.jl:178-197
    z, P = oracle.rouwenhorst(0.95, 0.0, 0.35, 5)
    assert np.allclose(P.sum(1), 1.0, rtol=1e-14)
    nu = math.sqrt(4 / (1 - 0.95 ** 2)) * 0.35
    assert z[0] == pytest.approx(-nu) and z[-1] == pytest.approx(nu) and z[2] == pytest.approx(0.0, abs=1e-15)
    q = (0.95 + 1) / 2
    assert P[0, 0] == pytest.approx(q ** 4) and P[0, 4] == pytest.approx((1 - q) ** 4)
    # stationary distribution of the Rouwenhorst chain is Binomial(n-1, 1/2)
    pi = np.array([math.comb(4, k) for k in range(5)]) / 16
    assert np.allclose(pi @ P, pi, atol=1e-14)


def test_scalegrid(oracle):  # utils.jl:200-238
    g = oracle.scalegrid(0.0, 50.0, 100, 2)
    assert g[0] == pytest.approx(0.0, abs=1e-14) and g[-1] == pytest.approx(50.0, rel=1e-14)
    assert np.all(np.diff(g) > 0) and np.all(np.diff(g, 2) > 0)
    t = np.linspace(math.log(math.log(1.0) + 1), math.log(math.log(51.0) + 1), 100)
    assert np.allclose(g, np.exp(np.exp(t) - 1) - 1, rtol=1e-13, atol=1e-14)
    gneg = oracle.scalegrid(-2.0, 10.0, 20, 1)
    assert gneg[0] == pytest.approx(-2.0) and gneg[-1] == pytest.approx(10.0)


def test_linrange_matches_julia_range(oracle):
    # Julia: collect(range(0.0, stop=50.0, length=500))[[2, 3, 250, 499]]
    g = oracle.linrange(0.0, 50.0, 500)
    from fractions import Fraction
    for i in (1, 2, 249, 333, 498):
        assert g[i] == float(Fraction(50 * i, 499))
    assert g[0] == 0.0 and g[-1] == 50.0


def test_minimal_variants_oracle():
    """minimal_EGM_borrow / minimal_EGM_bequest (minimal_EGM.jl:49-159), unpinned by reference tests:
    structural checks of the restatement."""
    from dcegm_jl_b200 import Param, _abi
    from oracle import oracle

    p = Param(par=dict(na=40))
    a = oracle.solve_minimal([p])
    b = oracle.solve_minimal_borrow([p])            # a_low = 0: identical to minimal_EGM
    assert np.array_equal(a.lines[0, :, 0, 0, 0], b.lines[0, :, 0, 0, 0])
    pb = Param(par=dict(na=40, a_low=-2.0))
    b2 = oracle.solve_minimal_borrow([pb])
    x, y = b2.line(0, _abi.KIND_C, 0, 3)
    assert x[0] == -2.0 and y[0] == 0.0 and len(x) == 41 and np.all(np.diff(x) > 0)
    q = Param(par=dict(na=40, nu=0.0))
    c = oracle.solve_minimal_bequest([q])
    xc, yc = c.line(0, _abi.KIND_C, 0, q.nT - 1)
    xv, yv = c.line(0, _abi.KIND_V, 0, q.nT - 1)
    assert len(xc) == 40 and np.array_equal(xc, yc) and np.all(yv == 0.0)   # minimal_EGM.jl:102-104
    xc, yc = c.line(0, _abi.KIND_C, 0, 0)
    xv, yv = c.line(0, _abi.KIND_V, 0, 0)
    assert np.array_equal(xc, xv) and xc[0] == 0.0 and yc[0] == 0.0 and np.all(np.diff(xc) > 0)
    q2 = Param(par=dict(na=40, nu=2.0))
    d = oracle.solve_minimal_bequest([q2])
    # with a bequest motive the rich save for it in the penultimate period (minimal_EGM.jl:124-133); the
    # luxury shifter bbar makes the poor consume more than without
    c_no = c.line(0, _abi.KIND_C, 0, q.nT - 2)[1]
    c_beq = d.line(0, _abi.KIND_C, 0, q.nT - 2)[1]
    assert np.all(c_beq[-5:] < c_no[-5:]) and c_beq[1] > c_no[1]
    assert d.status[0] == 0


def test_oracle_on_a_fine_grid_keeps_its_invariants(oracle):
    """BASELINE configs[3] geometry on a short horizon (the GPU side is tests/test_gpu_parity_f_big.py): the layout grows
    with na, sections are not capped, every stored line is strictly increasing in x while the status word is clean."""
    from dcegm_jl_b200 import Param

    d = dict(na=4000, ny=8, nT=8, retage=9, beta=0.95, R=1.05, sigma=0.45)
    d["lambda"] = 5.45e-5
    r = oracle.solve_f([Param(par=d)], portable=True)
    lay = r.lay
    assert lay.na == 4000 and lay.cap == (4000 + 400 + 2 * 250 + 2 + 7) // 8 * 8 and lay.isec_cap == 250
    assert r.status[0] == 0 and r.lens.max() <= lay.cap
    assert r.meta[0, :, 0, 1].sum() > 0  # kinks were found
    for it in range(7):
        for idd in (0, 1):
            for kind in (0, 1):
                n = r.lens[0, it, 0, idd, kind]
                x = r.lines[0, it, 0, idd, kind, :n, 0]
                assert (np.diff(x) > 0).all()
