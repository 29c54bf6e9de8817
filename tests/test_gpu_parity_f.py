"""CUDA path (through the C ABI) vs the CPU oracle for the FModel retirement problem
(dc_EGM!(::FModel), do_secondary, secondary_envelope).

Two comparisons:
  * oracle in `portable` mode (same reproducible pow/exp/log as the kernels): everything must be
    IDENTICAL — line lengths, section / intersection / removed counts, removed index sets, status
    and every stored double, bit for bit;
  * oracle with glibc libm (the mode pinned to the reference's golden files): discrete structure
    identical, ordinates within 1e-10 relative (north_star) — except where the reference algorithm
    itself amplifies 1-ulp libm differences (tests/test_oracle_sensitivity.py documents these)."""
import numpy as np
import pytest

from conftest import bm_param, ftest_param
from parity import assert_bitwise, assert_parity, libm_sensitivity, line_errors

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def solver():
    from dcegm_jl_b200.solver import Solver

    s = Solver(0)
    yield s
    s.close()


def c3_batch(n, na=500, seed=25032020):
    """SURVEY.md §8(d) C3 draws: log10(lambda) ~ U(log10 2e-6, log10 0.2), sigma ~ U(0, 0.6)"""
    from dcegm_jl_b200 import Param

    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2))
        sig = rng.uniform(0, 0.6)
        d = dict(na=na, beta=0.95, R=1.05, sigma=float(sig))
        d["lambda"] = float(lam)
        out.append(Param(par=d))
    return out


def test_F_test_model_matches_oracle_and_golden(solver, oracle, golden):
    from dcegm_jl_b200 import _abi

    p = ftest_param()
    g = solver.solve(_abi.MODEL_F, [p], keep_dirty=True)
    assert_bitwise(g, oracle.solve_f([p], keep_dirty=True, portable=True))
    r = oracle.solve_f([p], keep_dirty=True)
    assert g.status[0] == 0
    assert_parity(g, r)
    # pre-envelope lines (runfdemo's vdirty/cdirty): only libm separates the two
    d = np.abs(g.dirty[:, :24] - r.dirty[:, :24]) / (1e-2 + np.abs(r.dirty[:, :24]))
    assert d.max() < 1e-11
    # and directly against the reference's own golden files (test/F_test.jl:47-69)
    for i in (1, 2):
        for it in range(p.nT, 0, -1):
            mpo, mvf = golden[f"policy_{i}_{it}"], golden[f"value_{i}_{it}"]
            cx, cy = g.line(0, 0, i - 1, it - 1)
            vx, vy = g.line(0, 1, i - 1, it - 1)
            assert len(cx) == mpo.shape[1] and len(vx) == mvf.shape[1]
            o = np.argsort(mpo[0], kind="stable")
            assert np.max(np.abs(cx - mpo[0, o])) < 1e-4 and np.max(np.abs(cy - mpo[1, o])) < 1e-4
            if it < p.nT:
                assert np.max(np.abs(vx - mvf[0])) < 1e-4 and np.max(np.abs(vy - mvf[1])) < 5e-4


def test_bm_model(solver, oracle):
    from dcegm_jl_b200 import _abi

    p = bm_param()
    g = solver.solve(_abi.MODEL_F, [p])
    assert_bitwise(g, oracle.solve_f([p], portable=True))
    assert_parity(g, oracle.solve_f([p]))


def test_c3_batch_parity(solver, oracle):
    from dcegm_jl_b200 import _abi

    ps = c3_batch(256)
    g = solver.solve(_abi.MODEL_F, ps, keep_dirty=True)
    assert_bitwise(g, oracle.solve_f(ps, nthreads=8, keep_dirty=True, portable=True))
    # against the libm oracle (the mode pinned to the golden files): discrete structure identical for EVERY model and
    # every ordinate within 1e-10 relative (north_star).  Measured (tools/libm_flip_share.py, 1024 C3 draws): no model
    # changes structure under the 1-ulp differences between glibc and dc_math.h, max relative deviation 9.2e-12.
    r = oracle.solve_f(ps, nthreads=8)
    same = (g.lens == r.lens).all(axis=(1, 2, 3, 4)) & (g.meta == r.meta).all(axis=(1, 2, 3))
    assert same.all(), f"{(~same).sum()} models differ in structure from the libm oracle"
    assert np.array_equal(g.removed, r.removed) and np.array_equal(g.status, r.status)
    share, worst_kink, worst_rest = libm_sensitivity(g, r)
    assert share == 1.0 and worst_kink <= 1e-10 and worst_rest <= 1e-10, (share, worst_kink, worst_rest)


@pytest.mark.parametrize("na", [20, 50, 100, 333, 1000])
def test_grid_sizes(solver, oracle, na):
    from dcegm_jl_b200 import _abi

    ps = [ftest_param(na=na), bm_param(na=na, sigma=0.0), bm_param(na=na, **{"lambda": 0.01})]
    g = solver.solve(_abi.MODEL_F, ps)
    assert_bitwise(g, oracle.solve_f(ps, portable=True))
    assert_parity(g, oracle.solve_f(ps))


def test_param_branches(solver, oracle):
    """delta>0 (return to work), nu>0 (bequest), gamma==1 (log utility), pension>0, ny != 5"""
    from dcegm_jl_b200 import Param, _abi

    ps = [Param(par=dict(delta=0.3)), Param(par=dict(nu=2.0)), Param(par=dict(gamma=1.0)),
          Param(par=dict(pension=0.4)), Param(par=dict(delta=0.1, nu=1.0, bbar=3.0)),
          Param(par=dict(alpha=0.1, beta=0.9))]
    g = solver.solve(_abi.MODEL_F, ps)
    assert_bitwise(g, oracle.solve_f(ps, portable=True))
    assert_parity(g, oracle.solve_f(ps))
    for ny, nT in ((3, 10), (7, 25), (9, 5), (1, 4)):
        ps = [Param(par=dict(ny=ny, nT=nT)), Param(par=dict(ny=ny, nT=nT, sigma=0.1))]
        g = solver.solve(_abi.MODEL_F, ps)
        assert_bitwise(g, oracle.solve_f(ps, portable=True))
        assert_parity(g, oracle.solve_f(ps))


def test_fetch_paths_agree(solver, oracle):
    from dcegm_jl_b200 import _abi

    ps = c3_batch(5, na=200, seed=3)
    g = solver.solve(_abi.MODEL_F, ps)
    # single-line fetch
    for (m, kind, id, it) in [(0, 0, 0, 3), (4, 1, 0, 20), (2, 1, 1, 0), (3, 0, 1, 24)]:
        x, y = solver.fetch_line(m, kind, id, it)
        gx, gy = g.line(m, kind, id, it)
        assert np.array_equal(x, gx) and np.array_equal(y, gy, equal_nan=True)
    # dense packing
    xy, off = solver.fetch_packed()
    assert off[-1] == g.lens.sum()
    flat_lens = g.lens.reshape(-1)
    assert np.array_equal(np.diff(off), flat_lens)
    s = 0
    for m in range(5):
        for it in (0, 7, 24):
            for id in (0, 1):
                for kind in (0, 1):
                    slot = ((m * 25 + it) * 2 + id) * 2 + kind
                    gx, gy = g.line(m, kind, id, it)
                    seg = xy[off[slot]:off[slot + 1]]
                    assert np.array_equal(seg[:, 0], gx) and np.array_equal(seg[:, 1], gy, equal_nan=True)
    # partial fetch
    part = solver.fetch(first=2, count=2)
    assert np.array_equal(part.lens, g.lens[2:4])
    assert line_errors(part, type(g)(part.lay, g.lines[2:4], g.lens[2:4], None, None, None, None, None, None)) == 0.0


def test_summaries(solver, oracle):
    from dcegm_jl_b200 import _abi

    ps = c3_batch(8, na=200, seed=11)
    g = solver.solve(_abi.MODEL_F, ps)
    sm = solver.summaries()
    assert np.array_equal(sm["status"], g.status)
    assert np.array_equal(sm["tot_isects"], g.meta[..., 1].sum(axis=(1, 2)))
    assert np.array_equal(sm["tot_removed"], g.meta[..., 2].sum(axis=(1, 2)))
    assert np.array_equal(sm["tot_points"], g.lens.sum(axis=(1, 2, 3, 4)))
    assert np.array_equal(sm["max_len"], g.lens[:, :24].max(axis=(1, 2, 3, 4)))
    for m in range(8):
        cy = g.line(m, 0, 0, 0)[1].sum() + g.line(m, 0, 1, 0)[1].sum()
        assert sm["checksum_c"][m] == pytest.approx(cy, rel=1e-12)
        cx, cyy = g.line(m, 0, 0, 0)
        probe = np.interp(0.3 * 50.0, cx, cyy)
        assert sm["c1_probe"][m, 4] == pytest.approx(probe, rel=1e-12)


def test_errors_are_codes_not_crashes(solver):
    from dcegm_jl_b200 import Param, _abi, _lib

    with pytest.raises(_lib.DcegmError) as e:
        solver.solve(_abi.MODEL_F, [Param(par=dict(na=100)), Param(par=dict(na=200))])
    assert e.value.code == _abi.E_INVALID and "share" in str(e.value)
    # grids too fine for shared memory go to the fine-grid solvers (tests/test_gpu_parity_f_big.py,
    # tests/test_gpu_parity_g_big.py); their polylines are indexed with 16 bits
    for model in (_abi.MODEL_F, _abi.MODEL_G):
        with pytest.raises(_lib.DcegmError) as e:
            solver.solve(model, [Param(par=dict(na=60000))])
        assert e.value.code == _abi.E_INVALID and "16 bits" in str(e.value)


@pytest.mark.parametrize("n,na", [(5, 120), (700, 40)])
def test_packed_pipelined_solve_equals_resident_solution(solver, n, na):
    """dcegm_solve_f_packed (chunked solve overlapped with packing + D2H) delivers bit for bit what
    dcegm_solve_f + dcegm_fetch_packed deliver."""
    from dcegm_jl_b200 import _abi

    ps = c3_batch(n, na=na, seed=7)
    solver.solve(_abi.MODEL_F, ps, fetch=False)
    xy0, off0 = solver.fetch_packed()
    st0 = solver.status()
    xy1, off1 = solver.solve_packed(_abi.MODEL_F, ps)
    st1 = solver.status()
    assert np.array_equal(off0, off1)
    assert xy0.shape == xy1.shape
    assert np.array_equal(np.ascontiguousarray(xy0).view(np.uint64), np.ascontiguousarray(xy1).view(np.uint64))
    assert np.array_equal(st0, st1)
