"""Known answers of test/Envtests.jl (fixtures from src/plotting.jl) for the oracle's envelope
algebra, both in the literal (reference complexity, bisection) and the fast variant."""
import numpy as np
import pytest

LIT = [False, True]


def tplot1():  # plotting.jl:785-793
    n = 15
    x1 = np.linspace(0, 10, n)
    x2 = np.linspace(-1, 9, n)
    return [(x1, x1.copy()), (x2, np.ones(n) * 5)]


@pytest.mark.parametrize("literal", LIT)
def test_upper_env_test1(oracle, literal):  # Envtests.jl:20-30
    e = oracle.upper_env(tplot1(), literal)
    x, y = e["env"][:, 0], e["env"][:, 1]
    assert np.all(np.diff(x) >= 0)
    assert np.all(y[x < 5.0] == 5.0)
    assert np.all(y[x > 5.0] == x[x > 5.0])
    # the crossing at x=5 is a grid point -> no intersection, 17 points
    assert len(e["isects"]) == 0 and len(x) == 17


@pytest.mark.parametrize("literal", LIT)
def test_three_lines(oracle, literal):  # Envtests.jl:88-103
    x1 = np.arange(-0.9, 2.7 + 1e-9, 0.3)
    x2 = np.arange(0.0, 1.0 + 1e-9, 0.1)
    x3 = np.arange(1.0, 2.9 + 1e-9, 0.45)
    x1 = np.array([-0.9 + 0.3 * i for i in range(13)])
    x2 = np.array([0.0 + 0.1 * i for i in range(11)])
    x3 = np.array([1.0 + 0.45 * i for i in range(5)])
    e = oracle.upper_env([(x1, x1.copy()), (x2, 2 * x2), (x3, 0.1 * x3 + 1.9)], literal)
    x, y = e["env"][:, 0], e["env"][:, 1]
    assert np.all(np.diff(x) >= 0)
    assert np.all(y[x < 0.0] == x1[x1 < 0.0])
    m = (x > 0.0) & (x < 1.0)
    assert np.all(y[m] == (2 * x2)[(x2 > 0.0) & (x2 < 1.0)])
    assert len(e["isects"]) == 1
    assert np.allclose(e["isects"][0], [1.9 / 0.9, 1.9 / 0.9], rtol=1.5e-8, atol=0)  # isapprox default
    assert len(x) == 20


@pytest.mark.parametrize("literal", LIT)
def test_decreasing(oracle, literal):  # Envtests.jl:104-115, plotting.jl:929-944
    n = 15
    x1 = np.linspace(0, 10, n)
    x2 = np.linspace(-1, 9, n)
    x = np.concatenate([x1, x2])
    y = np.concatenate([x1[::-1], np.ones(n) * 5])
    secs = oracle.split_line(x, y)
    assert len(secs) == 3
    e = oracle.secondary_envelope(x, y, literal)
    ex = e["env"][:, 0]
    assert np.all(np.diff(ex) >= 0)
    assert len(e["isects"]) == 0 and len(ex) == 18


def test_splitline_simple(oracle):  # Envtests.jl:118-126
    x = [1, 2, 3, 1.5, 2.1, 2.9]
    y = [1, 1.5, 1.7, 1.2, 1.8, 2.1]
    secs = oracle.split_line(x, y)
    assert [s[0] for s in secs] == [3, 2, 3]


@pytest.mark.parametrize("literal", LIT)
def test_splitline_zigzag(oracle, literal):  # Envtests.jl:127-142
    x = [1, 2, 3, 2.9, 2.5, 1.9, 1.8, 1.5, 2.1, 2.9]
    y = [1, 1.5, 1.7, 1.6, 1.55, 1.4, 1.3, 1.2, 1.8, 2.1]
    secs = oracle.split_line(x, y)
    assert len(secs) == 3
    assert [(s[1], s[2]) for s in secs] == [(1.0, 3.0), (1.5, 3.0), (1.5, 2.9)]
    assert [s[0] for s in secs] == [3, 6, 3]
    e = oracle.secondary_envelope(x, y, literal)
    assert len(e["isects"]) == 1
    assert e["n_segments"] == 3
    # SURVEY.md App. B.3
    assert np.allclose(e["env"], [[1, 1], [1.6, 1.3], [2.1, 1.8], [2.9, 2.1], [3, 1.7]], rtol=0, atol=1e-12)
    assert np.allclose(e["isects"][0], [1.6, 1.3], atol=1e-12)
    assert list(e["removed"] + 1) == [2, 4, 5, 6, 7, 8]


def test_envelope_constructor_semantics(oracle):  # Envtests.jl:4-16
    n = 15
    x1 = np.linspace(0, 10, n)
    e = oracle.secondary_envelope(x1, x1)
    assert len(e["env"]) == n and len(e["isects"]) == 0 and len(e["removed"]) == 0


def test_interp_exact_at_knots_and_linear_extrapolation(oracle):  # line.jl:220-254
    x = np.array([0.0, 1.0, 3.0, 7.0])
    y = np.array([1.0, -2.0, 5.0, 5.5])
    ox, oy = oracle.interp(x, y, x)
    assert np.array_equal(ox, x) and np.array_equal(oy, y)
    ox, oy = oracle.interp(x, y, [-1.0, 9.0, 2.0])
    assert np.allclose(oy, [4.0, 5.75, 1.5], rtol=1e-15)
    assert np.allclose(ox, [-1.0, 9.0, 2.0], rtol=1e-15)


def test_literal_equals_fast_on_random_zigzags(oracle):
    rng = np.random.default_rng(7)
    for trial in range(200):
        n = int(rng.integers(6, 60))
        x = np.cumsum(rng.uniform(0.2, 1.0, n))
        # fold back a few stretches to create falling sections
        for _ in range(int(rng.integers(1, 4))):
            a = int(rng.integers(1, n - 3))
            b = int(rng.integers(a + 1, min(n - 1, a + 8)))
            x[a:b + 1] = x[a:b + 1][::-1] - rng.uniform(0, 0.5)
        if len(np.unique(x)) != n:
            continue
        y = np.log(1 + np.abs(x)) + rng.normal(0, 0.2, n)
        a_ = oracle.secondary_envelope(x, y, literal=False)
        b_ = oracle.secondary_envelope(x, y, literal=True)
        if a_["status"] or b_["status"]:
            assert a_["status"] == b_["status"]
            continue
        assert a_["n_segments"] == b_["n_segments"]
        assert len(a_["env"]) == len(b_["env"]) and len(a_["isects"]) == len(b_["isects"])
        assert list(a_["removed"]) == list(b_["removed"])
        assert np.allclose(a_["env"], b_["env"], rtol=1e-12, atol=1e-12)
