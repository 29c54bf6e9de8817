"""The host-side mirror of DCEGM.jl's entry points (dcegm.jl_b200/api.py).  The GPU tests read like
test/F_test.jl and test/basics.jl of the reference."""
import numpy as np
import pytest

import dcegm_jl_b200 as D
from dcegm_jl_b200 import _abi


def test_containers_without_gpu():
    # line.jl:37 exact Point equality; Envtests.jl:4-16 Envelope defaults
    assert D.Point(1.0, 2.0) == D.Point(1.0, 2.0)
    assert not (D.Point(1.0, 2.0) == D.Point(1.0, 2.0 + 1e-16 * 4))
    e = D.Envelope(D.MLine([[0.0, 0.0], [1.0, 1.0]]))
    assert D.bound(e) == 0.0 and D.gets(e) == [] and D.getr(e) == []
    assert np.array_equal(D.getx(e), [0.0, 1.0]) and np.array_equal(D.gety(e), [0.0, 1.0])
    assert e.env[2] == D.Point(1.0, 1.0) and len(e.env) == 2


def test_no_cpu_fallback(monkeypatch):
    """without a CUDA device the entry points raise — nothing is computed on the host"""
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(Exception) as ei:
        D.runf(par=dict(na=50))
    assert "CUDA" in str(ei.value) or "libdcegm" in str(ei.value)


@pytest.mark.gpu
def test_runf_like_F_test(golden):
    """test/F_test.jl:34-69 through runf(): equal point counts, 1e-4 / 5e-4 against the MATLAB files"""
    par = dict(na=500, beta=1 / 1.05, sigma=0.35, R=1.05, delta=0.0)
    par["lambda"] = 0.000002
    m, p = D.runf(par=par)
    assert p.na == 500 and p.nT == 25
    for i in (1, 2):
        for it in range(p.nT, 0, -1):
            mc = golden[f"policy_{i}_{it}"]
            mv = golden[f"value_{i}_{it}"]
            o = np.argsort(mc[0], kind="stable")  # F_test.jl:55
            assert len(m.c[i, it].env) == mc.shape[1]            # F_test.jl:53
            assert len(m.v[i, it].env) == mv.shape[1]            # F_test.jl:54
            assert np.max(np.abs(D.getx(m.c[i, it]) - mc[0][o])) < 1e-4
            assert np.max(np.abs(D.gety(m.c[i, it]) - mc[1][o])) < 1e-4
            if it == p.nT:
                assert abs(D.getx(m.v[i, it])[0] - mv[0][0]) < 1e-4 and abs(D.gety(m.v[i, it])[0] - mv[1][0]) < 1e-4
            else:
                assert np.max(np.abs(D.getx(m.v[i, it]) - mv[0])) < 1e-4
                assert np.max(np.abs(D.gety(m.v[i, it]) - mv[1])) < 5e-4
    # envelope bookkeeping of the worker value function: period 22 is the bend-back case of SURVEY App. B.1
    e = m.v[1, 22]
    assert len(D.gets(e)) == 4 and len(D.getr(e)) == 18 and D.bound(e) == D.gety(e)[0]
    assert all(1 <= r <= 549 for r in D.getr(e))


@pytest.mark.gpu
def test_runfdemo_rung_minimal_bm(oracle):
    m, p = D.runfdemo(par=dict(na=100))
    assert len(m.vdirty[1, 3]) == 100 and len(m.cdirty[2, 24]) == 100
    with pytest.raises(KeyError):
        m.vdirty[1, 25]
    g, p = D.rung(par=dict(na=60, ny=3, nT=6))
    ref = oracle.solve_g([p], portable=True)
    for iy in (1, 2, 3):
        x, y = ref.line(0, _abi.KIND_V, 0, 1, iy - 1)
        assert np.array_equal(D.getx(g.v[1, iy, 2]), x) and np.array_equal(D.gety(g.v[1, iy, 2]), y)
    with pytest.raises(IndexError):
        g.v[1, 4, 2]
    mm, cc = D.minimal_EGM(D.Param())
    assert len(mm) == 25 and len(mm[0]) == 101 and len(mm[24]) == 2
    assert np.all(np.diff(mm[0]) > 0) and cc[0][0] == 0.0
    t = D.bm(samples=3, verbose=False)
    assert t["samples"] == 3 and t["min_ms"] > 0


@pytest.mark.gpu
def test_reference_exception_mapping():
    """a parameterisation on which the reference's interpolate() throws (a repeated knot: the envelope finds an
    intersection at a turning point shared by two sections) raises here too, unless strict=False"""
    par = dict(na=500, beta=0.95, R=1.05, sigma=0.03651394092735729)
    par["lambda"] = 0.0001016080125633262
    with pytest.raises(D.DCEGMError):
        D.runf(par=par)
    m, _ = D.runf(par=par, strict=False)
    assert int(m.solution.status[0]) == _abi.ST_DUP_KNOT
