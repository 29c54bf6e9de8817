"""Forward simulator (sim(::FModel), sim(::GModel), simulate.jl:61-203) on the GPU against the CPU
oracle, through the C ABI: with the oracle in its portable mode (same dc_math / dc_rng arithmetic) every
panel entry, per-period moment, retirement-age count and status word is bit-identical."""
import numpy as np
import pytest

from dcegm_jl_b200 import Param, _abi
from oracle import oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def solver():
    from dcegm_jl_b200.solver import Solver

    s = Solver(0)
    yield s
    s.close()


def batch(model, n, nsims, seed=11):
    rng = np.random.default_rng(seed)
    out = []
    for k in range(n):
        d = dict(na=80 if model == _abi.MODEL_F else 60, nsims=nsims, rseed=1000 + k % 3, sigma=float(rng.uniform(0.05, 0.5)))
        d["lambda"] = float(10 ** rng.uniform(-4, -0.5))
        if model == _abi.MODEL_G:
            d.update(rho=float(rng.uniform(0, 0.9)), pension=float(rng.uniform(0, 1)), beta=float(rng.uniform(0.8, 1.0)))
        elif k % 2:
            d["delta"] = 0.2  # retirees may return to work: exercises rshocks (simulate.jl:127)
        out.append(Param(par=d))
    return out


def bits(a):
    """bit patterns, all NaNs mapped to one pattern (F: inc of period 1 stays NaN, simulate.jl:18)"""
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = a.view(np.uint64).copy()
    b[np.isnan(a)] = 0x7ff8000000000000
    return b


@pytest.mark.parametrize("model", [_abi.MODEL_F, _abi.MODEL_G])
def test_sim_bitwise_vs_oracle(solver, model):
    ps = batch(model, 6, 2500)  # 2500 agents: three tiles, the last one ragged
    g = solver.solve(model, ps)
    solver.simulate(panel=True)
    mom, hist, st = solver.sim_moments()
    r = oracle.solve(model, ps, portable=True)
    assert np.array_equal(bits(g.lines), bits(r.lines))
    for m in (0, 3, 5):
        omom, ohist, ost, opan = oracle.simulate(model, ps, r, portable=True, panel_model=m)
        pan = solver.sim_panel(m, ps[m].nsims)
        for k in ("w0", "w1", "cons", "util", "inc", "prob_work"):
            assert np.array_equal(bits(pan[k]), bits(opan[k])), k
        for k in ("ystate", "worker", "ret_age"):
            assert np.array_equal(pan[k], opan[k]), k
        assert np.array_equal(bits(mom), bits(omom))
        assert np.array_equal(hist, ohist) and np.array_equal(st, ost)
    assert (hist.sum(axis=1) == 2500).all()


def test_sim_given_shocks_and_api(solver):
    """the reference's raw rand() matrices as input (what a Julia-side caller would pass)"""
    p = Param(par={"na": 70, "lambda": 0.02})
    rng = np.random.default_rng(5)
    ns = 300
    sh = dict(wshocks=rng.random((ns, p.nT)), dshocks=rng.random((ns, p.nT)), rshocks=rng.random((ns, p.nT)),
              w0shocks=rng.random(ns))
    g = solver.solve(_abi.MODEL_F, [p])
    solver.simulate(panel=True, shocks=sh)
    pan = solver.sim_panel(0, ns)
    r = oracle.solve(_abi.MODEL_F, [p], portable=True)
    _, _, _, opan = oracle.simulate(_abi.MODEL_F, [p], r, shocks=sh, portable=True, panel_model=0)
    for k in ("w0", "w1", "cons", "util", "prob_work"):
        assert np.array_equal(bits(pan[k]), bits(opan[k])), k
    assert np.array_equal(pan["worker"], opan["worker"])
    # libm-mode oracle: same discrete paths, values within 1e-10
    r2 = oracle.solve(_abi.MODEL_F, [p])
    _, _, _, lpan = oracle.simulate(_abi.MODEL_F, [p], r2, shocks=sh, panel_model=0)
    assert np.array_equal(pan["worker"], lpan["worker"])
    assert np.allclose(pan["cons"], lpan["cons"], rtol=1e-10, atol=0)


def test_sim_api_mirror():
    import dcegm_jl_b200 as D

    s = D.rs(par={"na": 60, "nsims": 200})
    assert s.cons.shape == (200, s.p.nT) and np.all(s.worker[:, 0] == 1)
    assert np.allclose(s.moments["cons"], s.cons.mean(axis=0), rtol=1e-12)
    sg = D.rsg(par={"na": 50, "nsims": 150})
    assert sg.ystate.min() >= 1 and sg.ystate.max() <= sg.p.ny


def test_sim_sweep_moments_only(solver):
    """sweep mode: many models, moments only, different nsims per model"""
    ps = batch(_abi.MODEL_F, 40, 100)
    for k, p in enumerate(ps):
        p.nsims = 50 + 37 * (k % 5)
    g = solver.solve(_abi.MODEL_F, ps)
    solver.simulate()
    mom, hist, st = solver.sim_moments()
    r = oracle.solve(_abi.MODEL_F, ps, portable=True)
    omom, ohist, ost, _ = oracle.simulate(_abi.MODEL_F, ps, r, portable=True, nthreads=4)
    assert np.array_equal(bits(mom), bits(omom)) and np.array_equal(hist, ohist) and np.array_equal(st, ost)
    assert np.array_equal(hist.sum(axis=1), [p.nsims for p in ps])
