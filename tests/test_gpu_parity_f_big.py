"""Fine-grid retirement solver (csrc/solver_f_big.cu: working lines in global memory, the EGM step spread
over the whole GPU, two launches per period) through the C ABI vs the CPU oracle and vs the shared-memory
resident solver.  BASELINE.json configs[3] ("fine-grid retirement model: 20000 asset points, 50 periods,
32 quadrature nodes") is the case it exists for.  The bar is the same as for solve_f_kernel: against the
oracle's portable mode every stored double, length, section / intersection / removed count, removed index
set, status word and summary is bit-identical; against the libm oracle the discrete structure is exact and
ordinates agree within 1e-10 relative."""
import numpy as np
import pytest

from conftest import bm_param, ftest_param
from parity import assert_bitwise, assert_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def solver():
    from dcegm_jl_b200.solver import Solver

    s = Solver(0)
    yield s
    s.close()


def draws(n, seed=4, **kw):
    """SURVEY.md §8(d) C3/C4 law: log10(lambda) ~ U(log10 2e-6, log10 0.2), sigma ~ U(0, 0.6)"""
    from dcegm_jl_b200 import Param

    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        d = dict(beta=0.95, R=1.05, sigma=float(rng.uniform(0, 0.6)), **kw)
        d["lambda"] = float(10 ** rng.uniform(np.log10(2e-6), np.log10(0.2)))
        out.append(Param(par=d))
    return out


def solve_both(solver, ps, **kw):
    """(resident solver, fine-grid solver) solutions + summaries; asserts they are identical"""
    from dcegm_jl_b200 import _abi

    a = solver.solve(_abi.MODEL_F, ps, **kw)
    sa = solver.summaries()
    b = solver.solve(_abi.MODEL_F, ps, fine_grid=True, **kw)
    sb = solver.summaries()
    assert_bitwise(b, a)
    assert sa.tobytes() == sb.tobytes()
    return a, b


def test_forced_fine_grid_equals_resident_solver(solver, oracle):
    """the two CUDA solvers on the F_test / bm models and a C3 draw: identical bits, dirty lines included"""
    from dcegm_jl_b200 import _abi

    for p in (ftest_param(), bm_param()):
        a, b = solve_both(solver, [p], keep_dirty=True)
        assert_bitwise(b, oracle.solve_f([p], keep_dirty=True, portable=True))
    solve_both(solver, draws(96, na=500))


def test_branches_bequest_and_return_to_work(solver, oracle):
    """nu > 0 (bequest period, generic node path) and delta > 0 on the fine-grid solver"""
    from dcegm_jl_b200 import Param, _abi

    ps = []
    for nu, delta, lam in ((2.0, 0.0, 1e-3), (0.0, 0.3, 2e-6), (1.0, 0.2, 0.05)):
        d = dict(na=300, nu=nu, delta=delta)
        d["lambda"] = lam
        ps.append(Param(par=d))
    a, g = solve_both(solver, ps)
    assert_bitwise(g, oracle.solve_f(ps, portable=True))


@pytest.mark.parametrize("na,ny,nT,n", [(2000, 8, 10, 12), (5000, 16, 8, 4)])
def test_grids_beyond_shared_memory(solver, oracle, na, ny, nT, n):
    """automatic selection (the lines of one model no longer fit into 227 KB)"""
    from dcegm_jl_b200 import _abi

    ps = draws(n, na=na, ny=ny, nT=nT, retage=nT + 1)
    g = solver.solve(_abi.MODEL_F, ps)
    assert g.lay.cap >= 2048
    assert_bitwise(g, oracle.solve_f(ps, portable=True))
    assert_parity(g, oracle.solve_f(ps))


def test_c4_geometry_short_horizon(solver, oracle):
    """BASELINE configs[3] geometry (na = 20000, ny = 32) on a short horizon so that the oracle finishes in seconds"""
    from dcegm_jl_b200 import Param, _abi

    ps = []
    for lam, sig in ((2e-6, 0.35), (1e-2, 0.1)):
        d = dict(na=20000, ny=32, nT=6, retage=51, beta=0.95, R=1.05, sigma=sig)
        d["lambda"] = lam
        ps.append(Param(par=d))
    g = solver.solve(_abi.MODEL_F, ps)
    r = oracle.solve_f(ps, portable=True)
    assert_bitwise(g, r)
    assert g.meta[..., 1].sum() > 0  # the envelope found kinks


def test_more_sections_than_the_resident_solver_holds(solver, oracle):
    """at na = 20000 splitLine returns hundreds of sections (223 in period 5 here; the shared-memory solver stops
    at 192): the fine-grid solver has room for one section per point"""
    from dcegm_jl_b200 import Param, _abi

    d = dict(na=20000, ny=32, nT=12, retage=51, beta=0.95, R=1.05, sigma=0.45)
    d["lambda"] = 5.45e-5
    ps = [Param(par=d)]
    g = solver.solve(_abi.MODEL_F, ps)
    r = oracle.solve_f(ps, portable=True)
    assert r.meta[..., 0].max() > 192 and r.status[0] == 0
    assert_bitwise(g, r)


def test_c4_full_size_model(solver, oracle):
    """one model of BASELINE configs[3] at its full size (na = 20000, ny = 32, nT = 50, retage = 51; the first
    (lambda, sigma) draw of the C4 batch of tools/config_runs.py) against the oracle (~30 s of CPU), plus the
    size-independent properties of a second one: every stored line strictly increasing in x unless the status word
    says otherwise, lengths within the layout, the summary consistent with the lines"""
    from dcegm_jl_b200 import Param, _abi

    rng = np.random.default_rng(25032020)
    lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), 256)
    sig = rng.uniform(0, 0.6, 256)
    ps = []
    for i in (0, 1):
        d = dict(na=20000, ny=32, nT=50, retage=51, beta=0.95, R=1.05, sigma=float(sig[i]))
        d["lambda"] = float(lam[i])
        ps.append(Param(par=d))
    g = solver.solve(_abi.MODEL_F, ps)
    sm = solver.summaries()
    r = oracle.solve_f(ps[:1], portable=True)
    assert r.meta[..., 0].max() > 192  # more sections than the shared-memory solver could hold
    for name in ("lens", "meta", "removed", "vbound", "isects", "status"):
        assert np.array_equal(getattr(g, name)[:1], getattr(r, name)), name
    k = np.arange(r.lay.cap)
    m = np.broadcast_to((k[None, None, None, None, None, :] < r.lens[..., None])[..., None], r.lines.shape)
    a, b = g.lines[:1][m], r.lines[m]
    assert ((a == b) | (np.isnan(a) & np.isnan(b))).all()
    for mi in (0, 1):
        assert g.lens[mi].max() <= g.lay.cap and sm["max_len"][mi] == g.lens[mi, :49].max()
        assert sm["tot_segments"][mi] == g.meta[mi, :, 0, 0].sum() and sm["tot_isects"][mi] == g.meta[mi, :, 0, 1].sum()
        if g.status[mi] == 0:
            for it in range(49):
                for idd in (0, 1):
                    for kind in (0, 1):
                        x = g.line(mi, kind, idd, it)[0]
                        assert (np.diff(x) > 0).all()


def test_packed_transfer_of_fine_grid_solution(solver):
    """dcegm_solve_f_packed routes through the same solver selection"""
    from dcegm_jl_b200 import _abi

    ps = draws(3, na=2000, ny=8, nT=6, retage=7)
    g = solver.solve(_abi.MODEL_F, ps)
    xy, off = solver.solve_packed(_abi.MODEL_F, ps)
    lens = g.lens.reshape(-1)
    assert np.array_equal(np.diff(off), np.minimum(lens, g.lay.cap))
    flat = g.lines.reshape(-1, g.lay.cap, 2)
    for s in (0, 5, len(lens) - 1):
        assert np.array_equal(xy[off[s]:off[s + 1]], flat[s, : lens[s]], equal_nan=True)


def test_irregular_models_on_the_fine_grid(solver, oracle):
    """a wider draw at na = 3000: models whose lines stop being strictly increasing (status NONMONOTONE / MISALIGNED, the
    reference would throw there) take the sequential insertion path, the merge sort of the policy and the generic EGM node;
    they must still match the oracle bit for bit, status words included"""
    from dcegm_jl_b200 import _abi

    ps = draws(32, seed=11, na=3000, ny=8, nT=14, retage=15)
    g = solver.solve(_abi.MODEL_F, ps)
    r = oracle.solve_f(ps, portable=True, nthreads=8)
    assert_bitwise(g, r)
    assert (r.status != 0).any(), "the draw should contain irregular models"
