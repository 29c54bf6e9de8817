"""CPU checks of the simulator oracle (oracle_simulate) and of the shared shock source (dc_rng.h):
Philox4x32-10 known-answer vectors, the normal quantile against scipy, and the accounting identities
of simulate.jl:61-203 on a solved model.  The reference has no test of sim(): parity unpinned."""
import ctypes as C

import numpy as np
import pytest

from dcegm_jl_b200 import Param, _abi
from oracle import oracle


def test_philox_known_answers():
    # Random123 kat_vectors: philox4x32-10
    lib = oracle.lib()
    out = (C.c_uint32 * 4)()
    lib.oracle_philox(0, 0, 0, 0, 0, 0, out)
    assert [hex(v) for v in out] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    lib.oracle_philox(0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, out)
    assert [hex(v) for v in out] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    lib.oracle_philox(0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344, 0xa4093822, 0x299f31d0, out)
    assert [hex(v) for v in out] == ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


def test_uniform_and_normal_quantile():
    from scipy.stats import norm

    lib = oracle.lib()
    lib.oracle_sim_uniform.restype = C.c_double
    lib.oracle_norm_quantile.restype = C.c_double
    lib.oracle_norm_quantile.argtypes = [C.c_double]
    u = np.array([lib.oracle_sim_uniform(C.c_int64(25032020), i, 3, 1) for i in range(20000)])
    assert 0.0 <= u.min() and u.max() < 1.0
    assert abs(u.mean() - 0.5) < 0.01 and abs(u.var() - 1 / 12) < 0.005
    ps = np.concatenate([np.linspace(1e-12, 1 - 1e-12, 2001), [1e-300, 0.075, 0.925, 0.5]])
    q = np.array([lib.oracle_norm_quantile(float(x)) for x in ps])
    ref = norm.ppf(ps)
    assert np.max(np.abs(q - ref) / np.maximum(1.0, np.abs(ref))) < 5e-15
    assert lib.oracle_norm_quantile(0.0) == -np.inf and lib.oracle_norm_quantile(1.0) == np.inf


@pytest.mark.parametrize("model", [_abi.MODEL_F, _abi.MODEL_G])
def test_simulation_identities(model):
    par = dict(na=60, nsims=1500)
    if model == _abi.MODEL_F:
        par["lambda"] = 0.05
    p = Param(par=par)
    sol = oracle.solve(model, [p])
    mom, hist, st, pan = oracle.simulate(model, [p], sol, panel_model=0)
    assert st[0] == 0
    nT, ns = p.nT, p.nsims
    w0, w1, cons, inc, worker = pan["w0"], pan["w1"], pan["cons"], pan["inc"], pan["worker"]
    assert np.all((w0[:, 0] >= p.initw0) & (w0[:, 0] <= p.initw1))         # simulate.jl:84
    assert np.all(worker[:, 0] == 1)
    assert np.allclose(w0[:, 1:], w1[:, :-1] * p.R + inc[:, 1:], rtol=0, atol=1e-12)  # simulate.jl:103
    free = w0 - cons
    if model == _abi.MODEL_F:
        free = np.clip(free, p.a_low, p.a_high)                            # simulate.jl:109
    assert np.array_equal(w1, free)
    # retirement is absorbing with delta = 0 and ret_age marks the first retired period (simulate.jl:95)
    for i in range(0, ns, 97):
        ret = np.flatnonzero(worker[i] == 2)
        if len(ret):
            assert np.all(worker[i, ret[0]:] == 2) and pan["ret_age"][i] == ret[0] + 1
        else:
            assert pan["ret_age"][i] == 0
    assert hist[0].sum() == ns and np.array_equal(hist[0], np.bincount(pan["ret_age"], minlength=nT + 1))
    assert np.all(pan["prob_work"][worker == 2] == 0.0)
    # moments are means of the panel columns (the summation order differs from numpy's: tolerance)
    for j, k in enumerate(("w0", "w1", "cons", "util")):
        assert np.allclose(mom[0, :, j], pan[k].mean(axis=0), rtol=1e-12)
    assert np.allclose(mom[0, :, 6], (worker == 1).mean(axis=0))
    if model == _abi.MODEL_G:
        assert pan["ystate"].min() >= 1 and pan["ystate"].max() <= p.ny


def test_given_shocks_override_the_generator():
    p = Param(par={"na": 50, "nsims": 64, "lambda": 0.1})
    sol = oracle.solve(_abi.MODEL_F, [p])
    rng = np.random.default_rng(3)
    sh = dict(wshocks=rng.random((64, p.nT)), dshocks=rng.random((64, p.nT)), rshocks=rng.random((64, p.nT)),
              w0shocks=rng.random(64))
    _, _, _, a = oracle.simulate(_abi.MODEL_F, [p], sol, shocks=sh, panel_model=0)
    sh["dshocks"] = np.ones((64, p.nT))  # nobody ever stays a worker
    _, _, _, b = oracle.simulate(_abi.MODEL_F, [p], sol, shocks=sh, panel_model=0)
    assert np.array_equal(a["w0"][:, 0], p.initw0 + sh["w0shocks"] * (p.initw1 - p.initw0))
    assert np.all(b["worker"][:, 1:] == 2) and np.all(b["ret_age"] == 2)
