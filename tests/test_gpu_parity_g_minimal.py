"""CUDA path vs oracle for dc_EGM!(::GModel) (dc_algo.jl:112-206) and minimal_EGM (minimal_EGM.jl:3-43).
The reference has no test for either: parity here is GPU == oracle (bitwise in the
oracle's portable-math mode, 1e-10 against the libm oracle)."""
import numpy as np
import pytest

from parity import assert_bitwise, assert_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def solver():
    from dcegm_jl_b200.solver import Solver

    s = Solver(0)
    yield s
    s.close()


def g_draws(n, seed=25032020, nu=False):
    """SURVEY.md §8(d) C5 ranges (interact.jl:106-116,129)"""
    from dcegm_jl_b200 import Param

    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        d = dict(beta=float(rng.uniform(0.5, 1)), R=float(rng.uniform(1, 1.5)), alpha=float(rng.uniform(0, 0.5)),
                 sigma=float(rng.uniform(0.01, 0.35)), rho=float(rng.uniform(0, 0.95)), pension=float(rng.uniform(0, 1)))
        d["lambda"] = float(10 ** rng.uniform(np.log10(2e-7), 0))
        if nu:
            d["nu"] = float(rng.uniform(0, 10))
            d["bbar"] = float(rng.uniform(0.1, 10))
        out.append(Param(par=d))
    return out


def test_gmodel_defaults_and_variants(solver, oracle):
    from dcegm_jl_b200 import Param, _abi

    ps = [Param(), Param(par=dict(rho=0.0)), Param(par=dict(rho=0.5, nu=2.0)), Param(par=dict(nu=5.0, bbar=3.0)),
          Param(par=dict(gamma=1.0)), Param(par=dict(pension=0.5, rho=0.3))]
    g = solver.solve(_abi.MODEL_G, ps, keep_dirty=True)
    assert g.lay.nstate == 5
    assert_bitwise(g, oracle.solve_g(ps, keep_dirty=True, portable=True))
    assert_parity(g, oracle.solve_g(ps), rtol=1e-9)


def test_gmodel_c5_draws(solver, oracle):
    from dcegm_jl_b200 import _abi

    for nu in (False, True):
        ps = g_draws(192, nu=nu)
        g = solver.solve(_abi.MODEL_G, ps)
        assert_bitwise(g, oracle.solve_g(ps, nthreads=8, portable=True))


# ny against the kernel's thread groups (four income states per round, two when ny <= 2): part of a round, exactly
# one, one and a part, exactly two, two and one
@pytest.mark.parametrize("na,ny,nT", [(50, 3, 8), (200, 7, 12), (100, 1, 5), (60, 2, 6), (80, 4, 6), (60, 8, 5), (120, 9, 4)])
def test_gmodel_sizes(solver, oracle, na, ny, nT):
    from dcegm_jl_b200 import Param, _abi

    ps = [Param(par=dict(na=na, ny=ny, nT=nT)), Param(par=dict(na=na, ny=ny, nT=nT, rho=0.0, sigma=0.2))]
    g = solver.solve(_abi.MODEL_G, ps)
    assert_bitwise(g, oracle.solve_g(ps, portable=True))


def test_gmodel_two_and_four_thread_groups_agree(solver, oracle):
    """DCEGM_G_GROUPS forces the number of thread groups of solve_g_kernel (read at launch): both give the oracle's bits"""
    import os

    from dcegm_jl_b200 import _abi

    ps = g_draws(64)
    ref = oracle.solve_g(ps, nthreads=8, portable=True)
    try:
        for groups in ("2", "4"):
            os.environ["DCEGM_G_GROUPS"] = groups
            assert_bitwise(solver.solve(_abi.MODEL_G, ps), ref)
    finally:
        os.environ.pop("DCEGM_G_GROUPS", None)


def test_minimal_egm(solver, oracle):
    from dcegm_jl_b200 import Param, _abi

    rng = np.random.default_rng(1)
    ps = [Param()] + [Param(par=dict(beta=float(rng.uniform(0.5, 1.0)), R=float(rng.uniform(1.0, 1.5)))) for _ in range(63)]
    g = solver.solve(_abi.MODEL_MINIMAL, ps)
    r = oracle.solve_minimal(ps)
    assert np.array_equal(g.status, r.status) and np.array_equal(g.lens, r.lens)
    # no transcendental function on this path: bitwise against the plain oracle too
    for m in range(len(ps)):
        for it in range(25):
            gx, gy = g.line(m, 0, 0, it)
            rx, ry = r.line(m, 0, 0, it)
            assert np.array_equal(gx, rx) and np.array_equal(gy, ry)
    assert g.length(0, 0, 0, 24) == 2 and g.length(0, 0, 0, 0) == 101


def test_minimal_egm_sizes(solver, oracle):
    from dcegm_jl_b200 import Param, _abi

    ps = [Param(par=dict(na=500, ny=9, nT=40, sigma=0.2)), Param(par=dict(na=500, ny=9, nT=40, beta=0.99))]
    g = solver.solve(_abi.MODEL_MINIMAL, ps)
    r = oracle.solve_minimal(ps)
    assert np.array_equal(g.lens, r.lens)
    for m in range(2):
        for it in range(40):  # slots beyond len keep whatever an earlier solve left there
            n = g.length(m, 0, 0, it)
            assert np.array_equal(g.lines[m, it, 0, 0, 0, :n], r.lines[m, it, 0, 0, 0, :n])


@pytest.mark.parametrize("model", ["borrow", "bequest"])
def test_minimal_variants_bitwise(model):
    """minimal_EGM_borrow / minimal_EGM_bequest (minimal_EGM.jl:49-159): GPU == oracle bit for bit
    (portable math for u(c) in the bequest value function)."""
    from dcegm_jl_b200 import Param, _abi
    from dcegm_jl_b200.solver import Solver
    from oracle import oracle

    rng = np.random.default_rng(3)
    ps = []
    for k in range(24):
        d = dict(na=100, beta=float(rng.uniform(0.5, 1.0)), R=float(rng.uniform(1.0, 1.5)), sigma=float(rng.uniform(0.05, 0.5)))
        if model == "borrow":
            d["a_low"] = float(-rng.uniform(0, 3))
        else:
            d["nu"] = float(rng.uniform(0, 3)) if k % 3 else 0.0
            d["gamma"] = 1.0 if k % 5 == 0 else float(rng.uniform(1.2, 3))
        ps.append(Param(par=d))
    mid = _abi.MODEL_MINIMAL_BORROW if model == "borrow" else _abi.MODEL_MINIMAL_BEQUEST
    with Solver(0) as s:
        g = s.solve(mid, ps)
    r = oracle.solve(mid, ps, portable=True)
    assert np.array_equal(g.lens, r.lens) and np.array_equal(g.status, r.status)
    assert np.array_equal(g.lines.view(np.uint64), r.lines.view(np.uint64))
    if model == "bequest":
        import dcegm_jl_b200 as D

        m, c, v = D.minimal_EGM_bequest(ps[1])
        assert len(m) == ps[1].nT and len(m[0]) == 101 and len(m[-1]) == 100 and np.isfinite(v[0]).all()
