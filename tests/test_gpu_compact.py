"""Compact line delivery (dcegm_solve_*_compact / dcegm_fetch_compact): expanding the compact records on the host
gives the densely packed points of dcegm_fetch_packed BIT FOR BIT — retirement model incl. envelope-heavy and
status-flagged draws, fine-grid solver, general model — in about a third of the bytes."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def c3_draws(n, na=500):
    from dcegm_jl_b200 import Param

    rng = np.random.default_rng(25032020)  # the first n of the 4096 C3 draws (bench.py)
    lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), size=4096)
    sig = rng.uniform(0, 0.6, size=4096)
    ps = []
    for i in range(n):
        d = dict(na=na, beta=0.95, R=1.05, sigma=float(sig[i]))
        d["lambda"] = float(lam[i])
        ps.append(Param(par=d))
    return ps


def check(s, cl, ratio_max):
    xy, off = s.fetch_packed()
    exy, eoff = cl.expand(nthreads=4)
    assert np.array_equal(off, eoff)
    assert exy.tobytes() == xy.tobytes()
    full = xy.nbytes + off.nbytes
    assert cl.nbytes < ratio_max * full, (cl.nbytes, full)
    return cl.nbytes / full


def test_retirement_batch_pipeline_equals_packed():
    from dcegm_jl_b200 import _abi
    from dcegm_jl_b200.solver import Solver

    ps = c3_draws(700)  # three chunks of the pipeline; draw 91 is status-flagged (duplicate knot)
    with Solver(0) as s:
        cl = s.solve_compact(_abi.MODEL_F, ps)
        assert int(s.status()[91]) == _abi.ST_DUP_KNOT
        r = check(s, cl, 0.36)
        # the non-pipelined encoder of the resident solution gives the same stream
        cl2 = s.fetch_compact()
        assert cl2.words.tobytes() == cl.words.tobytes() and np.array_equal(cl2.woffsets, cl.woffsets)
        assert np.array_equal(cl2.aux, cl.aux) and np.array_equal(cl2.lens, cl.lens)
        # a sub-range, and single lines through the handle-free decoder
        cl3 = s.fetch_compact(first=90, count=3)
        sol = s.fetch(first=90, count=3)
        for it in (0, 16, 21, 24):
            for id_ in (0, 1):
                for kind in (_abi.KIND_C, _abi.KIND_V):
                    x, y = cl3.line(1, kind, id_, it)
                    xs, ys = sol.line(1, kind, id_, it)
                    assert x.tobytes() == xs.tobytes() and y.tobytes() == ys.tobytes()
    print(f"compact / packed bytes: {r:.3f}")


def test_fine_grid_and_general_model():
    from dcegm_jl_b200 import Param, _abi
    from dcegm_jl_b200.solver import Solver

    with Solver(0) as s:
        s.solve(_abi.MODEL_F, c3_draws(6, na=2400), fetch=False)  # fine-grid solver (lines in global memory)
        check(s, s.fetch_compact(), 0.36)
        rng = np.random.default_rng(3)
        pg = []
        for _ in range(40):
            d = dict(na=100, beta=float(rng.uniform(0.5, 1)), R=float(rng.uniform(1, 1.5)), sigma=float(rng.uniform(0, 0.35)),
                     rho=float(rng.uniform(0, 0.95)), alpha=float(rng.uniform(0, 0.5)))
            d["lambda"] = float(10 ** rng.uniform(np.log10(2e-7), 0))
            pg.append(Param(par=d))
        cl = s.solve_compact(_abi.MODEL_G, pg)
        check(s, cl, 0.50)
        # bequest variant: terminal lines on the full grid
        s.solve(_abi.MODEL_G, [Param(par=dict(na=100, nu=2.0))], fetch=False)
        check(s, s.fetch_compact(), 0.55)


def test_minimal_models_and_small_batches():
    """slots without a value line (minimal EGM stores (m, c) only), batches below one chunk, a batch of one"""
    from dcegm_jl_b200 import Param, _abi
    from dcegm_jl_b200.solver import Solver

    with Solver(0) as s:
        s.solve(_abi.MODEL_MINIMAL, [Param(par=dict(beta=b)) for b in (0.9, 0.95, 0.99)], fetch=False)
        check(s, s.fetch_compact(), 0.80)  # 101-point lines: the per-slot words weigh more
        cl = s.solve_compact(_abi.MODEL_F, c3_draws(1, na=60))
        check(s, cl, 0.60)
        cl = s.solve_compact(_abi.MODEL_F, c3_draws(300, na=60), keep_dirty=True)  # 300 models: two chunks, one launch
        check(s, cl, 0.60)
        with pytest.raises(Exception):
            s.solve(_abi.MODEL_F, c3_draws(2, na=60), summary_only=True)
            s.fetch_compact()
