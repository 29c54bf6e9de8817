"""How much does the *reference algorithm* move when pow/exp/log change by <= 1 ulp?
The oracle is run twice — glibc libm vs the reproducible dc_math.h functions — on the same models.
This is the floor for any comparison between two implementations with different math libraries
(including Julia vs MATLAB, or Julia on two machines): SURVEY.md App. B.4."""
import numpy as np

from conftest import ftest_param
from parity import RTOL, libm_sensitivity


def test_golden_model_structure_is_insensitive(oracle, golden):
    a = oracle.solve_f([ftest_param()])
    b = oracle.solve_f([ftest_param()], portable=True)
    assert np.array_equal(a.lens, b.lens) and np.array_equal(a.meta, b.meta) and np.array_equal(a.removed, b.removed)
    share, worst_kink, worst_rest = libm_sensitivity(b, a)
    assert share == 1.0 and worst_rest <= RTOL and worst_kink <= RTOL
    # the portable-math oracle reproduces the golden files just as well (F_test.jl tolerances)
    for i in (1, 2):
        for it in range(1, 25):
            mpo, mvf = golden[f"policy_{i}_{it}"], golden[f"value_{i}_{it}"]
            cx, cy = b.line(0, 0, i - 1, it - 1)
            vx, vy = b.line(0, 1, i - 1, it - 1)
            o = np.argsort(mpo[0], kind="stable")
            assert len(cx) == mpo.shape[1] and len(vx) == mvf.shape[1]
            assert np.max(np.abs(cx - mpo[0, o])) < 1e-4 and np.max(np.abs(cy - mpo[1, o])) < 1e-4
            assert np.max(np.abs(vx - mvf[0])) < 1e-4 and np.max(np.abs(vy - mvf[1])) < 5e-4


def test_c3_draws(oracle):
    from dcegm_jl_b200 import Param

    rng = np.random.default_rng(25032020)
    ps = []
    for _ in range(48):
        d = dict(na=300, beta=0.95, R=1.05, sigma=float(rng.uniform(0, 0.6)))
        d["lambda"] = float(10 ** rng.uniform(np.log10(2e-6), np.log10(0.2)))
        ps.append(Param(par=d))
    a = oracle.solve_f(ps, nthreads=8)
    b = oracle.solve_f(ps, nthreads=8, portable=True)
    same = (a.lens == b.lens).all(axis=(1, 2, 3, 4)) & (a.meta == b.meta).all(axis=(1, 2, 3))
    assert same.mean() >= 0.95
    keep = np.nonzero(same)[0]
    import copy

    from dcegm_jl_b200.solution import Solution

    def sub(s):
        lay = copy.copy(s.lay)
        lay.nmodels = len(keep)
        return Solution(lay, s.lines[keep], s.lens[keep], s.vbound[keep], s.meta[keep], s.isects[keep],
                        s.removed[keep], None, s.status[keep])
    share, worst_kink, worst_rest = libm_sensitivity(sub(b), sub(a))
    assert share >= 0.999 and worst_rest < 1e-6
