"""Host decoder of the compact line format (include/dcegm.h "compact lines", csrc/compact_host.cpp) — no GPU needed:
records built by the numpy restatement of the format's rules (dcegm_jl_b200.compact.encode_reference) from real
polylines of the CPU oracle decode back to the same bits; inconsistent records are rejected."""
import ctypes as C

import numpy as np
import pytest

from dcegm_jl_b200 import Param, _abi, _lib
from dcegm_jl_b200.compact import KIND_EXPLICIT, KIND_GRID, KIND_REF, encode_reference

_dp = C.POINTER(C.c_double)


def decode(grid, rc, nc, ac, rv, nv, av):
    lib = _lib.load()
    xy_c = np.zeros((max(nc, 1), 2))
    xy_v = np.zeros((max(nv, 1), 2))
    grid = np.ascontiguousarray(grid, float)
    rc = np.ascontiguousarray(rc, float)
    rv = np.ascontiguousarray(rv, float)
    code = lib.dcegm_decode_compact_pair(grid.ctypes.data_as(_dp), len(grid), rc.ctypes.data_as(_dp), nc, ac,
                                         rv.ctypes.data_as(_dp), nv, av, xy_c.ctypes.data_as(_dp), xy_v.ctypes.data_as(_dp))
    return code, xy_c[:nc], xy_v[:nv]


def test_oracle_lines_roundtrip(oracle):
    # small lambda: sections, intersections, kink pairs, bend-back points
    ps = []
    for lam, sig in ((2e-6, 0.35), (1e-3, 0.1)):
        d = dict(na=120, sigma=sig)
        d["lambda"] = lam
        ps.append(Param(par=d))
    sol = oracle.solve_f(ps)
    grid = oracle.linrange(ps[0].a_low, ps[0].a_high, 120)  # avec = range(a_low, a_high, na), param.jl:199
    words_compact = words_full = 0
    for m in range(2):
        for it in range(sol.lay.nT):
            for id_ in range(2):
                nc, nv = int(sol.lens[m, it, 0, id_, 0]), int(sol.lens[m, it, 0, id_, 1])
                Lc = sol.lines[m, it, 0, id_, 0, :nc]
                Lv = sol.lines[m, it, 0, id_, 1, :nv]
                rc, ac, rv, av = encode_reference(Lc, Lv, grid)
                code, xc, xv = decode(grid, rc, nc, ac, rv, nv, av)
                assert code == 0
                assert xc.tobytes() == np.ascontiguousarray(Lc).tobytes()
                assert xv.tobytes() == np.ascontiguousarray(Lv).tobytes()
                words_compact += len(rc) + len(rv)
                words_full += 2 * (nc + nv)
    assert words_compact < 0.42 * words_full  # x dropped, ordinates packed: about a third of the bytes at na = 120


def test_run_kinds_and_rejection():
    grid = np.array([0.0, 1.0, 2.0, 3.0])
    # policy: (0,0) grid 0; (1.5,0.5) grid 1; (7.25, 0.125) explicit; (3.5, 0.5) grid 3
    Lc = np.array([[0.0, 0.0], [1.5, 0.5], [7.25, 0.125], [3.5, 0.5]])
    Lv = np.array([[0.0, -1.0], [1.5, -0.5], [2.0, 9.0], [3.5, 0.25]])
    rc, ac, rv, av = encode_reference(Lc, Lv, grid)
    assert (ac & 0xFFFF, ac >> 16) == (3, 1) and (av & 0xFFFF, av >> 16) == (3, 1)
    ywc, ywv = int(rc[:1].view(np.uint64)[0]), int(rv[:1].view(np.uint64)[0])  # head word: size of the ordinate section
    # I_0, I_1 - I_0, one word of control bytes, one block of 4 residuals (at most 8 bytes each)
    assert 2 + 1 < ywc <= 2 + 1 + 4 and 2 + 1 < ywv <= 2 + 1 + 4
    runs_c = rc[1 + ywc:1 + ywc + 3].view(np.uint64)
    assert [(int(v & 0xFFFF), int(v >> 16 & 0xFFFF), int(v >> 32)) for v in runs_c] == [(0, 0, KIND_GRID), (2, 0, KIND_EXPLICIT), (3, 3, KIND_GRID)]
    runs_v = rv[1 + ywv:1 + ywv + 3].view(np.uint64)
    assert [(int(v & 0xFFFF), int(v >> 16 & 0xFFFF), int(v >> 32)) for v in runs_v] == [(0, 0, KIND_REF), (2, 0, KIND_EXPLICIT), (3, 3, KIND_REF)]
    code, xc, xv = decode(grid, rc, 4, ac, rv, 4, av)
    assert code == 0 and np.array_equal(xc, Lc) and np.array_equal(xv, Lv)
    # a run that points beyond the grid, a missing explicit word, a run list that does not start at 0
    r0 = 1 + ywc  # first run word
    bad = rc.copy()
    bad[r0:r0 + 1] = np.array([0 | (3 << 16) | (KIND_GRID << 32)], np.uint64).view(np.float64)
    assert decode(grid, bad, 4, ac, rv, 4, av)[0] == _abi.E_INVALID
    assert decode(grid, rc, 4, 3, rv, 4, av)[0] == _abi.E_INVALID
    bad = rc.copy()
    bad[r0:r0 + 1] = np.array([1 | (0 << 16) | (KIND_GRID << 32)], np.uint64).view(np.float64)
    assert decode(grid, bad, 4, ac, rv, 4, av)[0] == _abi.E_INVALID
    # an ordinate section that claims more words than the line can need, a residual width above 16 four-bit units
    bad = rc.copy()
    bad[0:1] = np.array([1000], np.uint64).view(np.float64)
    assert decode(grid, bad, 4, ac, rv, 4, av)[0] == _abi.E_INVALID
    bad = rc.copy()
    wb = bad[3:4].view(np.uint8)
    wb[0] = 17
    assert decode(grid, bad, 4, ac, rv, 4, av)[0] == _abi.E_INVALID
    wb[0] = 0x40 | 4  # unknown control bit
    assert decode(grid, bad, 4, ac, rv, 4, av)[0] == _abi.E_INVALID
    # empty slots
    assert decode(grid, np.zeros(1), 0, 0, np.zeros(1), 0, 0)[0] == 0


def test_ordinate_packing_roundtrip_on_hostile_values():
    """differences of the bit patterns wrap modulo 2^64: any doubles survive, whatever their order of magnitude"""
    rng = np.random.default_rng(5)
    grid = np.array([0.0, 1.0])
    for n in (1, 2, 3, 4, 7, 8, 9, 10, 15, 16, 17, 33, 34, 35, 70, 300):
        y = rng.standard_normal(n) * 10.0 ** rng.integers(-300, 300, n)
        y[rng.integers(0, n, 2)] = [np.inf, -0.0]
        if n > 5:
            y[5] = np.nan
            y[n // 2:n // 2 + 3] = 1.25  # a stretch of equal values: zero residuals
        L = np.stack([np.arange(n) * 7.0 + 100.0, y], axis=1)
        rc, ac, rv, av = encode_reference(L, L[:0], grid)
        code, xc, _ = decode(grid, rc, n, ac, np.zeros(1), 0, 0)
        assert code == 0 and xc.tobytes() == L.tobytes(), n
    smooth = np.stack([np.linspace(1, 2, 500) + 100, np.sqrt(np.linspace(1, 2, 500))], axis=1)
    rc, ac, _, _ = encode_reference(smooth, smooth[:0], grid)
    yw = int(rc[:1].view(np.uint64)[0])
    assert yw <= 0.5 * 500  # smooth ordinates: under 4 bytes per value instead of 8
    # every residual width 0 .. 16 and both difference orders occur and decode: y_k built from chosen residuals
    for order3 in (False, True):
        for w in range(0, 17):
            n = 29
            I = [0x3FF0000000000000, 0x3FF0000000000000 + 12345]
            d1, d2 = I[1] - I[0], 0
            for k in range(2, n):
                mag = 0 if w == 0 else (1 << (4 * w - 2)) + k  # needs exactly w four-bit units with its sign
                r = -mag if k & 1 else mag
                d2 = d2 + r if order3 else r
                d1 += d2
                I.append((I[-1] + d1) & ((1 << 64) - 1))
            y = np.array(I, np.uint64).view(np.float64)
            L = np.stack([np.arange(n) * 3.0 + 50.0, y], axis=1)
            rc, ac, _, _ = encode_reference(L, L[:0], grid)
            code, xc, _ = decode(grid, rc, n, ac, np.zeros(1), 0, 0)
            assert code == 0 and xc.tobytes() == L.tobytes(), (order3, w)


def test_negative_zero_and_nan_abscissae_survive():
    grid = np.array([0.0, 1.0])
    Lc = np.array([[-0.0, 0.0], [np.nan, 1.0], [np.inf, 2.0]])
    Lv = np.array([[-0.0, 1.0], [5.0, 2.0]])
    rc, ac, rv, av = encode_reference(Lc, Lv, grid)
    code, xc, xv = decode(grid, rc, 3, ac, rv, 2, av)
    assert code == 0 and xc.tobytes() == Lc.tobytes() and xv.tobytes() == Lv.tobytes()
