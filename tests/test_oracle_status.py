"""Status words of the CPU oracle follow the reference's own failure modes (include/dcegm.h):
  * a repeated abscissa produced by upper_env at a turning point shared by two sections -> DCEGM_ST_DUP_KNOT;
  * a defect that only touches period 1 (never interpolated by the solve, dc_algo.jl:22-37) is informational;
  * insert_right longer than insert_left is not an error (the loop of dc_algo.jl:486-492 runs over insert_left)."""
import numpy as np

from dcegm_jl_b200 import Param, _abi


def _p(**d):
    lam = d.pop("lam")
    d["lambda"] = lam
    return Param(par=d)


def test_duplicate_knot_from_shared_turning_point(oracle):
    # C3 draw no. 91 (tools/flag_census.py): period 17, the envelope holds (6.806393232369432, 4.580542787253808) twice
    p = _p(na=500, beta=0.95, R=1.05, sigma=0.05276548761156301, lam=0.00011833940640038141)
    for portable in (True, False):
        s = oracle.solve_f([p], portable=portable)
        assert int(s.status[0]) == _abi.ST_DUP_KNOT
        x, y = s.line(0, _abi.KIND_V, 0, 16)
        k = np.nonzero(np.diff(x) == 0)[0]
        assert k.size == 1 and y[k[0]] == y[k[0] + 1]           # the same Point twice ...
        isx = s.isect_points(0, 16)
        assert np.any((isx[:, 0] == x[k[0]]) & (isx[:, 1] == y[k[0]]))  # ... once as an intersection (envelope.jl:212)
        assert np.all(np.diff(x) >= 0)                          # never decreasing: issorted() holds
    assert _abi.ST_DUP_KNOT & _abi.ST_FATAL_MASK


def test_info_bits_are_not_fatal():
    assert _abi.ST_INFO_MASK == _abi.ST_EXTRA_RIGHT | _abi.ST_NONMONO_FIRST
    assert not (_abi.ST_INFO_MASK & _abi.ST_FATAL_MASK)
    from dcegm_jl_b200 import api

    api._raise_for_status(_abi.ST_EXTRA_RIGHT | _abi.ST_NONMONO_FIRST)  # returns normally, like runf


def test_first_period_defect_is_informational(oracle):
    """a model whose ONLY defect is a repeated knot in period 1: the reference's runf returns normally"""
    import sys, os

    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    from flag_census import draw

    model, arr = draw("c5", 2048)
    # model 1023 of the C5 draw carries only DCEGM_ST_NONMONO_FIRST (tools/flag_census.py c5)
    from dcegm_jl_b200.param import params_array

    found = None
    for c0 in range(0, 2048, 256):
        s = oracle.solve(model, params_array([arr[i] for i in range(c0, c0 + 256)]), nthreads=os.cpu_count() or 1)
        only = np.nonzero(s.status == _abi.ST_NONMONO_FIRST)[0]
        if only.size:
            found = (s, int(only[0]))
            break
    assert found is not None
    s, m = found
    lay = s.lay
    bad_periods = set()
    for it in range(lay.nT - 1):
        for iy in range(lay.nstate):
            for id_ in (0, 1):
                for kind in (0, 1):
                    n = int(s.lens[m, it, iy, id_, kind])
                    if not np.all(np.diff(s.lines[m, it, iy, id_, kind, :n, 0]) > 0):
                        bad_periods.add(it)
    assert bad_periods == {0}
