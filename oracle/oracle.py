"""ctypes front end of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY (see dcegm_oracle.cpp).

May be imported from tests/, __graft_entry__.smoke() and the cpu_baseline / --impl reference
legs of bench.py; never from dcegm.jl_b200/."""
import ctypes as C
import os
import subprocess

import numpy as np

import dcegm_jl_b200 as pkg
from dcegm_jl_b200 import _abi
from dcegm_jl_b200.param import params_array
from dcegm_jl_b200.solution import Solution

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "dcegm_oracle.cpp")
    hdr = os.path.join(_HERE, "..", "include", "dcegm.h")
    mth = os.path.join(_HERE, "..", "dcegm.jl_b200", "csrc", "dc_math.h")
    rng = os.path.join(_HERE, "..", "dcegm.jl_b200", "csrc", "dc_rng.h")
    srcs = [f for f in (src, hdr, mth, rng) if os.path.exists(f)]
    if force or not os.path.exists(so) or (srcs and os.path.getmtime(so) < max(os.path.getmtime(f) for f in srcs)):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        dp = C.POINTER(C.c_double)
        ip = C.POINTER(C.c_int)
        _LIB.oracle_norminv.restype = C.c_double
        _LIB.oracle_norminv.argtypes = [C.c_double]
        for f in ("oracle_u",):
            getattr(_LIB, f).restype = C.c_double
            getattr(_LIB, f).argtypes = [C.c_double, C.c_int, C.POINTER(_abi.dcegm_param)]
        for f in ("oracle_up", "oracle_iup"):
            getattr(_LIB, f).restype = C.c_double
            getattr(_LIB, f).argtypes = [C.c_double, C.POINTER(_abi.dcegm_param)]
        _LIB.oracle_income.restype = C.c_double
        _LIB.oracle_income.argtypes = [C.c_int, C.c_double, C.POINTER(_abi.dcegm_param)]
        _LIB.oracle_solve.restype = C.c_int
    return _LIB


def _p(a, t=C.c_double):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


def layout(model, nmodels, na, ny, nT, keep_dirty=False):
    lay = _abi.dcegm_layout()
    lib().oracle_layout(model, nmodels, na, ny, nT, int(keep_dirty), C.byref(lay))
    return lay


def solve(model, params, keep_dirty=False, literal=False, nthreads=1, portable=False):
    """Solve a batch of Param on the host.  Returns a Solution in the layout of include/dcegm.h."""
    arr = params_array(params)
    n = len(params)
    lay = layout(model, n, arr[0].na, arr[0].ny, arr[0].nT, keep_dirty)
    s = Solution.allocate(lay, keep_dirty)
    rc = lib().oracle_solve(
        C.c_int(model), arr, C.c_int(n), C.c_int(int(keep_dirty)), C.c_int(int(literal)), C.c_int(int(portable)), C.c_int(nthreads),
        _p(s.lines), _p(s.lens, C.c_int32), _p(s.vbound), _p(s.meta, C.c_int32), _p(s.isects),
        _p(s.removed, C.c_uint32), _p(s.dirty), _p(s.status, C.c_int32))
    if rc != 0:
        raise ValueError("oracle_solve: all models of a batch must share na, ny, nT")
    return s


def solve_f(params, **kw):
    return solve(_abi.MODEL_F, params, **kw)


def solve_g(params, **kw):
    return solve(_abi.MODEL_G, params, **kw)


def solve_minimal(params, **kw):
    return solve(_abi.MODEL_MINIMAL, params, **kw)


def solve_minimal_borrow(params, **kw):
    return solve(_abi.MODEL_MINIMAL_BORROW, params, **kw)


def solve_minimal_bequest(params, **kw):
    return solve(_abi.MODEL_MINIMAL_BEQUEST, params, **kw)


def secondary_envelope(x, y, literal=False):
    x = np.ascontiguousarray(x, float)
    y = np.ascontiguousarray(y, float)
    n = len(x)
    cap = 2 * n + 64
    env = np.zeros((cap, 2))
    isec = np.zeros((cap, 2))
    rem = np.zeros(cap, np.int32)
    ne, ni, nr, ns = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    st = lib().oracle_secondary_envelope(_p(x), _p(y), n, int(literal), _p(env), cap, C.byref(ne), _p(isec), cap,
                                         C.byref(ni), _p(rem, C.c_int32), cap, C.byref(nr), C.byref(ns))
    return dict(env=env[: ne.value], isects=isec[: ni.value], removed=rem[: nr.value], n_segments=ns.value, status=st)


def dc_pow(x, y):
    x = np.ascontiguousarray(x, float)
    y = np.ascontiguousarray(np.broadcast_to(y, x.shape), float)
    out = np.zeros_like(x)
    lib().oracle_dc_pow(_p(x), _p(y), x.size, _p(out))
    return out


def dc_exp(x):
    x = np.ascontiguousarray(x, float)
    out = np.zeros_like(x)
    lib().oracle_dc_exp(_p(x), x.size, _p(out))
    return out


def dc_log(x):
    x = np.ascontiguousarray(x, float)
    out = np.zeros_like(x)
    lib().oracle_dc_log(_p(x), x.size, _p(out))
    return out


def upper_env(lines, literal=False):
    """lines: list of (x, y) arrays, one per section"""
    xs = np.concatenate([np.asarray(l[0], float) for l in lines])
    ys = np.concatenate([np.asarray(l[1], float) for l in lines])
    off = np.zeros(len(lines) + 1, np.int32)
    off[1:] = np.cumsum([len(l[0]) for l in lines])
    cap = 2 * len(xs) + 64
    env = np.zeros((cap, 2))
    isec = np.zeros((cap, 2))
    ne, ni = C.c_int(), C.c_int()
    st = lib().oracle_upper_env(_p(xs), _p(ys), _p(off, C.c_int32), len(lines), int(literal), _p(env), cap,
                                C.byref(ne), _p(isec), cap, C.byref(ni))
    return dict(env=env[: ne.value], isects=isec[: ni.value], status=st)


def split_line(x, y):
    x = np.ascontiguousarray(x, float)
    y = np.ascontiguousarray(y, float)
    cap = len(x) + 2
    ln = np.zeros(cap, np.int32)
    lo = np.zeros(cap)
    hi = np.zeros(cap)
    ns = lib().oracle_split_line(_p(x), _p(y), len(x), _p(ln, C.c_int32), _p(lo), _p(hi), cap)
    return [(int(ln[s]), float(lo[s]), float(hi[s])) for s in range(ns)]


def interp(x, y, q):
    x = np.ascontiguousarray(x, float)
    y = np.ascontiguousarray(y, float)
    q = np.ascontiguousarray(q, float)
    ox = np.zeros(len(q))
    oy = np.zeros(len(q))
    lib().oracle_interp(_p(x), _p(y), len(x), _p(q), len(q), _p(ox), _p(oy))
    return ox, oy


def ccp(v1, v2, lam):
    v1 = np.ascontiguousarray(v1, float)
    v2 = np.ascontiguousarray(v2, float)
    out = np.zeros(len(v1))
    lib().oracle_ccp(_p(v1), _p(v2), len(v1), C.c_double(lam), _p(out))
    return out


def logsum(v1, v2, lam):
    v1 = np.ascontiguousarray(v1, float)
    v2 = np.ascontiguousarray(v2, float)
    out = np.zeros(len(v1))
    lib().oracle_logsum(_p(v1), _p(v2), len(v1), C.c_double(lam), _p(out))
    return out


def quadpoints(n, lo=0.0, hi=1.0):
    x = np.zeros(n)
    w = np.zeros(n)
    lib().oracle_quadpoints(n, C.c_double(lo), C.c_double(hi), _p(x), _p(w))
    return x, w


def norminv(p):
    return lib().oracle_norminv(float(p))


def gausshermite(n):
    x = np.zeros(n)
    w = np.zeros(n)
    lib().oracle_gausshermite(n, _p(x), _p(w))
    return x, w


def rouwenhorst(rho, mu, sigma, n):
    z = np.zeros(n)
    P = np.zeros((n, n))
    lib().oracle_rouwenhorst(C.c_double(rho), C.c_double(mu), C.c_double(sigma), n, _p(z), _p(P))
    return z, P


def scalegrid(lb, ub, n, logorder=1):
    out = np.zeros(n)
    lib().oracle_scalegrid(C.c_double(lb), C.c_double(ub), n, logorder, _p(out))
    return out


def linrange(lo, hi, n):
    out = np.zeros(n)
    lib().oracle_linrange(C.c_double(lo), C.c_double(hi), n, _p(out))
    return out


def simulate(model, params, sol, shocks=None, portable=False, nthreads=1, panel_model=-1):
    """Forward simulation (simulate.jl:61-203) of the models in `sol` (a Solution in the shared layout).
    Returns (moments [n,nT,8], ret_hist [n,nT+1], status [n], panel dict or None)."""
    from dcegm_jl_b200 import _abi as A

    arr = params_array(params)
    n = len(arr)
    nT = arr[0].nT
    mom = np.zeros((n, nT, A.SIM_NMOM))
    hist = np.zeros((n, nT + 1), np.int32)
    st = np.zeros(n, np.int32)
    sp, keep = None, None
    nsims = arr[max(panel_model, 0)].nsims
    if shocks is not None:
        ss, keep = A.make_sim_shocks(shocks, nT)
        sp = C.byref(ss)
        nsims = ss.nsims
    pd = pi = pr = None
    if panel_model >= 0:
        pd = np.zeros((6, nT, nsims))
        pi = np.zeros((2, nT, nsims), np.int32)
        pr = np.zeros(nsims, np.int32)
    lay = sol.lay
    rc = lib().oracle_simulate(
        C.c_int(model), arr, C.c_int(n), C.c_int(int(portable)), C.byref(lay), _p(sol.lines), _p(sol.lens, C.c_int32),
        _p(sol.vbound), sp, C.c_int(nthreads), _p(mom), _p(hist, C.c_int32), _p(st, C.c_int32), C.c_int(panel_model),
        _p(pd) if pd is not None else None, _p(pi, C.c_int32) if pi is not None else None,
        _p(pr, C.c_int32) if pr is not None else None)
    assert rc == 0
    panel = None
    if panel_model >= 0:
        names = ("w0", "w1", "cons", "util", "inc", "prob_work")
        panel = {k: pd[j].T for j, k in enumerate(names)}  # (nsims, nT) views of column-major storage
        panel["ystate"], panel["worker"], panel["ret_age"] = pi[0].T, pi[1].T, pr
    return mom, hist, st, panel
