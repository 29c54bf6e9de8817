"""Pins the CPU oracle against the reference's golden files — the checks of test/F_test.jl:47-69."""
import numpy as np

from conftest import ftest_param

# point counts per period quoted in SURVEY.md §8c (period 1..25)
WORKER_POLICY = [501, 501, 501, 500, 500, 500, 499, 499, 499, 498, 498, 499, 500, 501, 501, 500, 500, 496, 497, 497,
                 492, 540, 499, 501, 2]
WORKER_VALUE = [501, 501, 501, 500, 500, 500, 499, 498, 498, 498, 498, 498, 498, 498, 498, 496, 496, 493, 492, 490,
                486, 536, 498, 501, 2]


def test_F_test_against_matlab_csv(oracle, golden):
    p = ftest_param()
    s = oracle.solve_f([p])
    assert s.status[0] == 0
    for i in (1, 2):
        for it in range(p.nT, 0, -1):
            mpo = golden[f"policy_{i}_{it}"]
            mvf = golden[f"value_{i}_{it}"]
            cx, cy = s.line(0, 0, i - 1, it - 1)
            vx, vy = s.line(0, 1, i - 1, it - 1)
            assert len(cx) == mpo.shape[1], (i, it)      # F_test.jl:53
            assert len(vx) == mvf.shape[1], (i, it)      # F_test.jl:54
            fedsort = np.argsort(mpo[0], kind="stable")  # F_test.jl:55
            assert np.max(np.abs(cx - mpo[0, fedsort])) < 1e-4
            assert np.max(np.abs(cy - mpo[1, fedsort])) < 1e-4
            if it == p.nT:
                assert abs(vx[0] - mvf[0, 0]) < 1e-4 and abs(vy[0] - mvf[1, 0]) < 1e-4
            else:
                assert np.max(np.abs(vx - mvf[0])) < 1e-4
                assert np.max(np.abs(vy - mvf[1])) < 5e-4
                # tighter than F_test: the CSVs carry 8 significant digits
                assert np.max(np.abs(vy - mvf[1]) / np.maximum(1.0, np.abs(mvf[1]))) < 1e-7
                assert np.max(np.abs(cy - mpo[1, fedsort]) / np.maximum(1.0, np.abs(cy))) < 1e-7


def test_point_counts(oracle):
    s = oracle.solve_f([ftest_param()])
    assert [s.length(0, 0, 0, it) for it in range(25)] == WORKER_POLICY
    assert [s.length(0, 1, 0, it) for it in range(25)] == WORKER_VALUE
    assert [s.length(0, 0, 1, it) for it in range(25)] == [501] * 24 + [2]
    assert [s.length(0, 1, 1, it) for it in range(25)] == [501] * 24 + [2]


def test_envelope_activity_matches_survey_probe(oracle):
    """it=22 bend-back, 549 input points, 9 segments, 4 isects, 18 removed"""
    s = oracle.solve_f([ftest_param()])
    nseg, nis, nrm, npre = s.meta[0, 21, 0]
    assert (nseg, nis, nrm, npre) == (9, 4, 18, 49)
    nseg, nis, nrm, npre = s.meta[0, 22, 0]
    assert (nseg, nis, nrm, npre) == (5, 1, 53, 49)
    nseg, nis, nrm, npre = s.meta[0, 19, 0]
    assert (nseg, nis, nrm) == (17, 7, 18)
    assert (s.meta[0, :3, 0, 2] == 0).all()
    assert len(s.removed_idx(0, 21)) == 18


def test_literal_and_fast_envelope_agree_on_full_model(oracle):
    """the O(n^2) bisection variant (reference complexity) and the O(n log n) closed-form variant"""
    p = ftest_param(na=200)
    a = oracle.solve_f([p], literal=False)
    b = oracle.solve_f([p], literal=True)
    assert (a.lens == b.lens).all() and (a.meta == b.meta).all() and (a.removed == b.removed).all()
    fin = np.isfinite(a.lines) & np.isfinite(b.lines)
    assert (np.isfinite(a.lines) == np.isfinite(b.lines)).all()
    err = np.abs(a.lines[fin] - b.lines[fin]) / np.maximum(1.0, np.abs(a.lines[fin]))
    assert err.max() < 1e-11


def test_terminal_period(oracle):
    p = ftest_param()
    s = oracle.solve_f([p])
    for id in (0, 1):
        cx, cy = s.line(0, 0, id, 24)
        assert list(cx) == [0.0, 50.0] and list(cy) == [0.0, 50.0]   # dc_algo.jl:17
        vx, vy = s.line(0, 1, id, 24)
        assert list(vx) == [0.0, 50.0] and vy[0] == 0.0 and np.isnan(vy[1])  # dc_algo.jl:19


def test_threads_give_identical_results(oracle):
    ps = [ftest_param(na=120, sigma=0.1 * k) for k in range(1, 5)]
    a = oracle.solve_f(ps, nthreads=1)
    b = oracle.solve_f(ps, nthreads=4)
    assert np.array_equal(a.lines, b.lines, equal_nan=True) and (a.lens == b.lens).all()
