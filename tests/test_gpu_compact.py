"""Compact line delivery (dcegm_solve_*_compact / dcegm_fetch_compact): expanding the compact records on the host
gives the densely packed points of dcegm_fetch_packed BIT FOR BIT — retirement model incl. envelope-heavy and
status-flagged draws, fine-grid solver, general model — in about a third of the bytes."""
import os

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


def aliased(cl):
    """slots without a record of their own: non-empty, and an earlier non-empty slot starts at the same offset"""
    idx = np.flatnonzero(cl.lens > 0)
    _, first = np.unique(cl.woffsets[:-1][idx], return_index=True)
    al = np.zeros(len(cl.lens), bool)
    al[idx] = True
    al[idx[first]] = False
    return al


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
        r = check(s, cl, 0.22)
        # identical lines travel once: every retiree pair (id 1) and the worker's terminal pair of a chunk's later models
        # point at the chunk's first model (lambda and sigma enter neither), and nothing else does
        al = aliased(cl)
        per = 25 * 4
        assert al.reshape(700, 25, 2, 2)[:, :24, 0, :].sum() == 0
        roots = np.flatnonzero(~al.reshape(700, per)[:, 2])  # models that carry their own retiree records
        assert roots[0] == 0 and 2 <= len(roots) <= 8, roots
        assert al.sum() == (700 - len(roots)) * 52
        # the non-pipelined encoder of the resident solution: one root (model 0), same lines
        cl2 = s.fetch_compact()
        check(s, cl2, 0.22)
        assert aliased(cl2).sum() == 699 * 52
        assert np.array_equal(cl2.aux, cl.aux) and np.array_equal(cl2.lens, cl.lens)
        # without the detection both encoders give the same stream, in slot order
        os.environ["DCEGM_COMPACT_ALIAS"] = "0"
        try:
            cl_na = s.solve_compact(_abi.MODEL_F, ps)
            check(s, cl_na, 0.36)
            cl2_na = s.fetch_compact()
            assert cl2_na.words.tobytes() == cl_na.words.tobytes() and np.array_equal(cl2_na.woffsets, cl_na.woffsets)
            assert aliased(cl_na).sum() == 0 and (np.diff(cl_na.woffsets) >= 0).all()
            assert np.array_equal(cl_na.aux, cl.aux) and np.array_equal(cl_na.lens, cl.lens)
        finally:
            os.environ.pop("DCEGM_COMPACT_ALIAS")
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
        # five identical general models: everything behind the first one is a reference to it
        cl = s.solve_compact(_abi.MODEL_G, [Param(par=dict(na=100, rho=0.5))] * 5)
        check(s, cl, 0.20)
        assert aliased(cl).reshape(5, -1)[1:].sum() == (cl.lens.reshape(5, -1)[1:] > 0).sum() and aliased(cl).reshape(5, -1)[0].sum() == 0
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


def test_device_records_equal_the_reference_encoder():
    """the device encoder writes exactly the records the numpy restatement of the format builds (word for word):
    abscissa runs, control bytes, residual payload, padding"""
    from dcegm_jl_b200 import Param, _abi
    from dcegm_jl_b200.compact import encode_reference
    from dcegm_jl_b200.solver import Solver

    ps = []
    for lam, sig in ((2e-6, 0.35), (1e-3, 0.1), (0.05, 0.5)):
        d = dict(na=120, sigma=sig)
        d["lambda"] = lam
        ps.append(Param(par=d))
    os.environ["DCEGM_COMPACT_ALIAS"] = "0"  # every slot with its own record
    try:
        with Solver(0) as s:
            sol = s.solve(_abi.MODEL_F, ps)
            assert (sol.status.astype(np.int64) & _abi.ST_FATAL_MASK == 0).all()
            cl = s.fetch_compact()
            checked = 0
            for m in range(len(ps)):
                grid = cl.grid(m)
                for it in range(sol.lay.nT):
                    for id_ in range(2):
                        nc, nv = int(sol.lens[m, it, 0, id_, 0]), int(sol.lens[m, it, 0, id_, 1])
                        rc, ac, rv, av = encode_reference(sol.lines[m, it, 0, id_, 0, :nc], sol.lines[m, it, 0, id_, 1, :nv], grid)
                        sc, sv = cl.slot(m, _abi.KIND_C, id_, it), cl.slot(m, _abi.KIND_V, id_, it)
                        for slot, rec, aux in ((sc, rc, ac), (sv, rv, av)):
                            w0 = int(cl.woffsets[slot])
                            assert int(cl.aux[slot]) == aux, (m, it, id_, slot)
                            assert cl.words[w0:w0 + len(rec)].tobytes() == rec.tobytes(), (m, it, id_, slot)
                            checked += len(rec)
            assert checked == int(cl.woffsets[-1])  # the records tile the stream: nothing between or behind them
    finally:
        os.environ.pop("DCEGM_COMPACT_ALIAS")
