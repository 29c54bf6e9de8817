"""CSV writer/reader in the format of matlab/polyline.m:472-484 and F_test's checks on top of it
(dcegm.jl_b200/io.py), and v_analytic (plotting.jl:3-42).  CPU: the model object wraps an oracle solution;
GPU: the same through runf()."""
import numpy as np
import pytest

import dcegm_jl_b200 as D
from dcegm_jl_b200 import _abi, io
from dcegm_jl_b200.api import FModel


def _ftest_model_cpu():
    from conftest import ftest_param
    from oracle import oracle

    p = ftest_param()
    return FModel(p, oracle.solve(_abi.MODEL_F, [p]), _abi.MODEL_F), p


def test_csv_roundtrip_and_f_test(tmp_path, golden):
    m, p = _ftest_model_cpu()
    assert io.f_test(m, golden) == []                      # F_test.jl:47-69 against the MATLAB files
    files = io.write_csv(m, str(tmp_path))
    assert len(files) == 100
    first = open(files[0]).read().split("\n")
    assert len(first) == 3 and first[0].startswith("0,")    # two rows, %.8g like dlmwrite(...,'precision',8)
    back = io.read_csv(str(tmp_path), p.nT)
    assert io.f_test(m, back) == []
    # the written files agree with the reference's own files to the printed precision
    for (name, i, it), a in back.items():
        g = golden[f"{name}_{i}_{it}"]
        assert a.shape == g.shape
        if name == "policy":
            a, g = a[:, np.argsort(a[0], kind="stable")], g[:, np.argsort(g[0], kind="stable")]
        ok = np.isfinite(g)
        assert np.allclose(a[ok], g[ok], rtol=0, atol=5e-4 if name == "value" else 1e-4)
    # a perturbed solution is caught
    m.solution.lines[0, 3, 0, 0, _abi.KIND_C, 5, 1] += 1e-3
    assert io.f_test(m, golden) == [(1, 4, "policy y")]


def test_scale_grid_and_v_analytic():
    g = io.scale_grid(0.0, 50.0, 7, 1)
    assert g[0] == 0.0 and abs(g[-1] - 50.0) < 1e-12 and np.all(np.diff(np.diff(g)) > 0)   # utils.jl:208-217
    g2 = io.scale_grid(-2.0, 5.0, 9, 2)
    assert abs(g2[0] + 2.0) < 1e-12 and abs(g2[-1] - 5.0) < 1e-12
    m, p = _ftest_model_cpu()
    va = D.v_analytic(m, p, 1, 10)
    vx = D.getx(m.v[1, 10])
    assert va.shape == (p.k + len(vx) - 1, 2)
    assert np.all(np.diff(va[:, 0]) >= 0)
    # continuous where the analytic part meets the envelope's second point (plotting.jl:14)
    assert abs(va[p.k - 1, 0] - vx[1]) < 1e-9 and abs(va[p.k - 1, 1] - D.gety(m.v[1, 10])[1]) < 1e-9
    vt = D.v_analytic(m, p, 2, p.nT)
    assert vt.shape == (p.na, 2)


@pytest.mark.gpu
def test_f_test_through_csv_gpu(tmp_path, golden):
    par = dict(na=500, beta=1 / 1.05, sigma=0.35, R=1.05, delta=0.0)
    par["lambda"] = 0.000002
    m, p = D.runf(par=par)
    assert io.f_test(m, golden) == []
    io.write_csv(m, str(tmp_path))
    assert io.f_test(m, io.read_csv(str(tmp_path), p.nT)) == []
