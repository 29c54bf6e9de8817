"""Accuracy of the reproducible pow/exp/log shared by the CUDA kernels and the oracle's portable mode
(dcegm.jl_b200/csrc/dc_math.h), measured in ulp against mpmath at 200 bits."""
import mpmath as mp
import numpy as np

mp.mp.prec = 200


def max_ulp(got, exact):
    worst = 0.0
    for g, ex in zip(got, exact):
        if ex == 0:
            assert g == 0
            continue
        worst = max(worst, float(abs(mp.mpf(float(g)) - ex) / mp.mpf(float(np.spacing(abs(float(ex)))))))
    return worst


def test_log(oracle):
    rng = np.random.default_rng(1)
    xs = np.concatenate([10 ** rng.uniform(-3, 3, 2000), rng.uniform(0.5, 2, 2000), 1 + rng.uniform(-1e-3, 1e-3, 300),
                         10 ** rng.uniform(-300, 300, 300), [5e-324, 1e-310, 1.0, 2.0, 0.5]])
    assert max_ulp(oracle.dc_log(xs), [mp.log(mp.mpf(float(x))) for x in xs]) < 0.6
    assert oracle.dc_log([1.0])[0] == 0.0 and oracle.dc_log([0.0])[0] == -np.inf and np.isnan(oracle.dc_log([-1.0])[0])


def test_exp(oracle):
    rng = np.random.default_rng(2)
    xs = np.concatenate([rng.uniform(-745, 709, 2000), rng.uniform(-2, 2, 2000), rng.uniform(-40, 0, 1000)])
    assert max_ulp(oracle.dc_exp(xs), [mp.exp(mp.mpf(float(x))) for x in xs]) < 0.7
    # the properties the ccp/logsum shortcut of the kernel relies on
    assert oracle.dc_exp([0.0])[0] == 1.0
    assert np.array_equal(oracle.dc_exp([-745.3, -746.0, -1e4]), [0.0, 0.0, 0.0])
    assert oracle.dc_exp([710.0])[0] == np.inf
    assert oracle.dc_exp([-744.5])[0] == 5e-324  # gradual underflow


def test_pow(oracle):
    rng = np.random.default_rng(3)
    xs = 10 ** rng.uniform(-3, 2.5, 6000)
    ys = np.concatenate([np.full(1500, -1.95), np.full(1500, -0.95), np.full(1500, -1 / 1.95), rng.uniform(-6, 6, 1500)])
    ex = [mp.power(mp.mpf(float(x)), mp.mpf(float(y))) for x, y in zip(xs, ys)]
    assert max_ulp(oracle.dc_pow(xs, ys), ex) < 0.7
    got = oracle.dc_pow([0.0, 0.0, 1.0, 2.0, np.inf, 4.0], [2.0, -1.0, 5.0, 0.0, -1.0, 0.5])
    assert np.array_equal(got, [0.0, np.inf, 1.0, 1.0, 0.0, 2.0])
    assert np.isnan(oracle.dc_pow([-1.0], [0.5])[0])


def test_linrange_dd_matches_long_double(oracle):
    from oracle import oracle as o
    import ctypes as C

    o.lib().oracle_dc_linrange_at.restype = C.c_double
    o.lib().oracle_dc_linrange_at.argtypes = [C.c_double, C.c_double, C.c_int, C.c_int]
    rng = np.random.default_rng(4)
    bad = 0
    for _ in range(50):
        lo, hi = sorted(rng.uniform(0, 60, 2))
        n = int(rng.integers(3, 60))
        ref = o.linrange(lo, hi, n)
        got = np.array([o.lib().oracle_dc_linrange_at(lo, hi, n, i) for i in range(n)])
        bad += int((got != ref).sum())
        assert np.max(np.abs(got - ref) / np.spacing(np.abs(ref))) <= 1.0
    assert bad <= 3  # x87 double rounding differs from double-double in rare half-way cases
