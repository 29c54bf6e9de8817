"""run_sweep on the GPU library: a sweep that mixes grid sizes runs group by group, the combined record (summary +
simulation moments) equals what the separate calls return, and flagged models carry NaN moments."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_mixed_geometry_sweep_matches_separate_solves():
    from dcegm_jl_b200 import Param, _abi, sweep
    from dcegm_jl_b200.solver import Solver

    rng = np.random.default_rng(7)
    params = []
    for i in range(24):
        na, nT = [(60, 25), (100, 25), (80, 12)][i % 3]
        d = dict(na=na, nT=nT, nsims=2000, beta=float(rng.uniform(0.8, 1)), sigma=float(rng.uniform(0.05, 0.35)),
                 rho=float(rng.uniform(0, 0.9)))
        d["lambda"] = float(10 ** rng.uniform(-5, -1))
        params.append(Param(par=d))
    summ, mom = sweep.run_sweep(_abi.MODEL_G, params, simulate=True)
    assert mom.shape == (24, 25, _abi.SIM_NMOM)
    with Solver(0) as s:
        for i, p in enumerate(params):
            s.solve(_abi.MODEL_G, [p], fetch=False)
            one = s.summaries()[0]
            assert one.tobytes() == summ[i].tobytes()
            s.simulate()
            m1, _, _ = s.sim_moments()
            fatal = int(one["status"]) & (_abi.ST_FATAL_MASK | _abi.ST_NONMONO_FIRST)
            if fatal:
                assert np.isnan(mom[i]).all()
            else:
                assert mom[i, : p.nT].tobytes() == m1[0].tobytes()  # bit for bit (util means can be NaN in both)
                assert np.isnan(mom[i, p.nT:]).all()


def test_flagged_model_moments_are_masked():
    from dcegm_jl_b200 import Param, _abi, sweep

    # C3 draw no. 91: the reference's envelope repeats a knot in period 17 (tests/test_oracle_status.py)
    d = dict(na=500, beta=0.95, R=1.05, sigma=0.05276548761156301, nsims=500)
    d["lambda"] = 0.00011833940640038141
    good = dict(d, sigma=0.35)
    summ, mom = sweep.run_sweep(_abi.MODEL_F, [Param(par=d), Param(par=good)], simulate=True)
    assert int(summ["status"][0]) == _abi.ST_DUP_KNOT and int(summ["status"][1]) == 0
    assert np.isnan(mom[0]).all() and np.isfinite(mom[1][:, :3]).all()  # wealth and consumption means of the clean model
