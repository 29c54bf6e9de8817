"""Fine-grid solver of the general model (csrc/solver_g_big.cu: dc_EGM!(::GModel), dc_algo.jl:112-206, on lines that
live in global memory) and the simulator variant that reads the lines in place (sim_kernel<.., GL>):
  * forced onto small grids it gives the bits of the shared-memory resident solver and of the oracle (portable mode),
    summaries included;
  * on a grid that does not fit one SM's shared memory (the reference has no grid-size limit) solve + simulation run
    and match the oracle bit for bit."""
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


def draws(n, na, ny=5, nT=25, seed=5, nu=False, nsims=100):
    from dcegm_jl_b200 import Param

    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        d = dict(na=na, ny=ny, nT=nT, nsims=nsims, beta=float(rng.uniform(0.6, 1)), R=float(rng.uniform(1, 1.3)),
                 alpha=float(rng.uniform(0, 0.5)), sigma=float(rng.uniform(0.01, 0.35)), rho=float(rng.uniform(0, 0.95)),
                 pension=float(rng.uniform(0, 1)))
        d["lambda"] = float(10 ** rng.uniform(np.log10(2e-7), 0))
        if nu:
            d["nu"] = float(rng.uniform(0.5, 10))
            d["bbar"] = float(rng.uniform(0.1, 10))
        out.append(Param(par=d))
    return out


@pytest.mark.parametrize("nu", [False, True])
def test_forced_fine_grid_equals_resident_solver_and_oracle(solver, oracle, nu):
    from dcegm_jl_b200 import _abi

    ps = draws(24, 100, nu=nu)
    a = solver.solve(_abi.MODEL_G, ps, keep_dirty=True)
    sa = solver.summaries()
    b = solver.solve(_abi.MODEL_G, ps, keep_dirty=True, fine_grid=True)
    sb = solver.summaries()
    assert_bitwise(b, a)
    assert sa.tobytes() == sb.tobytes()
    assert_bitwise(b, oracle.solve_g(ps, nthreads=8, keep_dirty=True, portable=True))


def test_other_geometries_forced(solver, oracle):
    from dcegm_jl_b200 import _abi

    for na, ny, nT in ((50, 3, 8), (200, 7, 12), (100, 1, 5)):
        ps = draws(3, na, ny, nT, seed=na)
        b = solver.solve(_abi.MODEL_G, ps, fine_grid=True)
        assert_bitwise(b, oracle.solve_g(ps, portable=True))


def test_grid_beyond_shared_memory_solve_and_simulate(solver, oracle):
    """na = 1500, ny = 5: 2 x 5 line pairs of ~1700 points do not fit 227 KB; the resident solver used to return
    DCEGM_E_INVALID here"""
    from dcegm_jl_b200 import _abi

    ps = draws(3, 1500, ny=5, nT=10, seed=9, nsims=3000)
    g = solver.solve(_abi.MODEL_G, ps)
    r = oracle.solve_g(ps, nthreads=8, portable=True)
    assert_bitwise(g, r)
    assert_parity(g, oracle.solve_g(ps, nthreads=8), rtol=1e-9)
    assert g.meta[..., 1].sum() > 0  # kinks were found
    solver.simulate(panel=True)
    mom, hist, st = solver.sim_moments()
    omom, ohist, ost, opan = oracle.simulate(_abi.MODEL_G, ps, r, portable=True, panel_model=1)
    assert mom.tobytes() == omom.tobytes() or np.array_equal(np.nan_to_num(mom, nan=-7.0), np.nan_to_num(omom, nan=-7.0))
    assert np.array_equal(hist, ohist) and np.array_equal(st, ost)
    pan = solver.sim_panel(1, 3000)
    for k in ("w0", "w1", "cons", "util", "prob_work"):
        assert np.array_equal(np.nan_to_num(pan[k], nan=-7.0), np.nan_to_num(opan[k], nan=-7.0)), k
    assert np.array_equal(pan["ystate"], opan["ystate"]) and np.array_equal(pan["ret_age"], opan["ret_age"])


def test_retirement_model_simulation_on_a_fine_grid(solver, oracle):
    """the simulator reads lines of 3000+ points in place (they do not fit into shared memory)"""
    from dcegm_jl_b200 import Param, _abi

    ps = []
    for lam, sig in ((2e-4, 0.3), (0.05, 0.2)):
        d = dict(na=3000, nT=12, nsims=2500, sigma=sig, rseed=77)
        d["lambda"] = lam
        ps.append(Param(par=d))
    g = solver.solve(_abi.MODEL_F, ps)
    r = oracle.solve_f(ps, nthreads=2, portable=True)
    assert_bitwise(g, r)
    solver.simulate()
    mom, hist, st = solver.sim_moments()
    omom, ohist, ost, _ = oracle.simulate(_abi.MODEL_F, ps, r, portable=True, panel_model=0)
    assert np.array_equal(np.nan_to_num(mom, nan=-7.0), np.nan_to_num(omom, nan=-7.0))
    assert np.array_equal(hist, ohist) and np.array_equal(st, ost)
