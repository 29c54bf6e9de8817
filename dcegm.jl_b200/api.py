"""Host-side mirror of DCEGM.jl's entry points for the hot path, over the C ABI of include/dcegm.h.

Same names, argument meaning and error behaviour as the reference (no Julia toolchain exists in
this image, so the host side is Python; the Julia shim a maintainer would write is in
INTEGRATION.md):

    runf(par=...)       -> (FModel, Param)     src/dc_algo.jl:335-341
    runfdemo(par=...)   -> (FModel, Param)     src/dc_algo.jl:342-348   (keeps vdirty / cdirty)
    rung(par=...)       -> (GModel, Param)     src/dc_algo.jl:356-362
    minimal_EGM(p)      -> (m, c)              src/minimal_EGM.jl:3-43  (data instead of the plot)
    bm()                -> timing dict         src/bench.jl:3-31        (without the MATLAB shell-out)
    Point, MLine, Envelope, getx, gety, gets, getr, bound   src/line.jl, src/envelope.jl

Indices are 1-based like the reference: ``m.v[id, it]``, ``m.c[id, it]`` (FModel) and
``m.v[id, iy, it]`` (GModel).  ``getr`` returns 1-based indices into the line handed to
``secondary_envelope`` (including bend-back points), like ``Envelope.removed``.

Nothing here computes a solution on the host: every call goes to libdcegm_cuda.so and raises
if the library or a CUDA device is missing.
"""
import time

import numpy as np

from . import _abi
from .param import Param
from .solver import Solver

__all__ = ["Point", "MLine", "Envelope", "FModel", "GModel", "runf", "runfdemo", "rung", "minimal_EGM", "bm",
           "getx", "gety", "gets", "getr", "bound", "DCEGMError", "dc_EGM", "Simulation", "sim", "rs", "rsg",
           "minimal_EGM_borrow", "minimal_EGM_bequest", "v_analytic"]


class DCEGMError(RuntimeError):
    """the exception the reference would have thrown for this model"""


class Point:  # line.jl:1-4
    __slots__ = ("x", "y")

    def __init__(self, x, y):
        self.x, self.y = float(x), float(y)

    def __eq__(self, o):  # line.jl:37: exact equality on both coordinates
        return isinstance(o, Point) and self.x == o.x and self.y == o.y

    def __repr__(self):
        return f"Point({self.x!r}, {self.y!r})"


class MLine:  # line.jl:62-87: a vector of Points; stored here as an (n, 2) array view
    def __init__(self, xy):
        self.xy = np.asarray(xy, dtype=np.float64).reshape(-1, 2)

    @property
    def v(self):
        return [Point(x, y) for x, y in self.xy]

    def __len__(self):
        return self.xy.shape[0]

    def __getitem__(self, i):  # 1-based like the reference
        x, y = self.xy[i - 1]
        return Point(x, y)

    def __repr__(self):
        return f"{len(self)}-point MLine"


def getx(l):  # line.jl:121-124
    return (l.env if isinstance(l, Envelope) else l).xy[:, 0].copy()


def gety(l):
    return (l.env if isinstance(l, Envelope) else l).xy[:, 1].copy()


class Envelope:  # envelope.jl:16-40
    def __init__(self, env, isects=None, removed=None, vbound=0.0):
        self.env = env if isinstance(env, MLine) else MLine(env)
        self._isects = np.zeros((0, 2)) if isects is None else np.asarray(isects, dtype=np.float64).reshape(-1, 2)
        self.removed = [] if removed is None else [int(i) for i in removed]
        self.vbound = float(vbound)

    @property
    def isects(self):
        return [Point(x, y) for x, y in self._isects]

    def __repr__(self):
        return f"{len(self.env)}-point Float64 Envelope"


def gets(e):  # envelope.jl:52-54
    return e.isects


def getr(e):
    return e.removed


def bound(e):
    return e.vbound


class _Indexed:
    """1-based [id, it] / [id, iy, it] view building Envelope objects lazily from the flat Solution"""

    def __init__(self, sol, kind, nstate_indexed):
        self.sol, self.kind, self.g = sol, kind, nstate_indexed

    def __getitem__(self, key):
        if self.g:
            id, iy, it = key
        else:
            (id, it), iy = key, 1
        s, k = self.sol, self.kind
        lay = s.lay
        if not (1 <= id <= 2 and 1 <= it <= lay.nT and 1 <= iy <= lay.nstate):
            raise IndexError(f"BoundsError: attempt to access {key}")  # Julia BoundsError
        if k in (_abi.KIND_CDIRTY, _abi.KIND_VDIRTY):
            if s.dirty is None or it == lay.nT:
                raise KeyError("dirty lines exist only after runfdemo, for it < nT")
            return MLine(s.dirty[0, it - 1, iy - 1, id - 1, k - 2])
        n = int(s.lens[0, it - 1, iy - 1, id - 1, k])
        env = MLine(s.lines[0, it - 1, iy - 1, id - 1, k, :n])
        vb = float(s.vbound[0, it - 1, iy - 1, id - 1])
        if k == _abi.KIND_V and id == 1:
            return Envelope(env, s.isect_points(0, it - 1, iy - 1), s.removed_idx(0, it - 1, iy - 1) + 1, vb)
        return Envelope(env, vbound=vb if k == _abi.KIND_V else 0.0)


_STATUS_ERRORS = {
    _abi.ST_NONMONOTONE: "knot-vectors must be unique and sorted in increasing order",  # Interpolations.jl, line.jl:236
    _abi.ST_DUP_KNOT: "knot-vectors must be unique and sorted in increasing order",    # duplicate knot (envelope.jl:212,248-252)
    _abi.ST_NO_BRACKET: "ArgumentError: The interval [a,b] is not a bracketing interval",  # Roots.fzero, envelope.jl:199
    _abi.ST_ALL_FALLING: "an upper envelope requires by definition at least 2 lines.",      # envelope.jl:149-151
    _abi.ST_MISALIGNED: "BoundsError: insert_right shorter than insert_left",               # dc_algo.jl:486-491
    _abi.ST_INSERT_RANGE: "BoundsError: kink insertion index out of range",                 # dc_algo.jl:472,489
    _abi.ST_NONFINITE: "non-finite value in a computed line",
    _abi.ST_ISECT_OVERFLOW: "library capacity: more intersections than isec_cap",
    _abi.ST_LINE_OVERFLOW: "library capacity: line outgrew its slot",
    _abi.ST_SEG_OVERFLOW: "library capacity: more sections than seg_cap",
    _abi.ST_LOOP_LIMIT: "upper_env: third-line loop did not terminate",
}


def _raise_for_status(st):
    st &= _abi.ST_FATAL_MASK  # informational bits: the reference returns normally (include/dcegm.h)
    if st:
        raise DCEGMError("; ".join(t for b, t in _STATUS_ERRORS.items() if st & b) + f" (status 0x{st:x})")


class _Model:
    def __init__(self, p, sol, model):
        g = model == _abi.MODEL_G
        self.p, self.solution = p, sol
        self.v = _Indexed(sol, _abi.KIND_V, g)
        self.c = _Indexed(sol, _abi.KIND_C, g)
        self.vdirty = _Indexed(sol, _abi.KIND_VDIRTY, g)
        self.cdirty = _Indexed(sol, _abi.KIND_CDIRTY, g)


class FModel(_Model):  # param.jl:171-224
    pass


class GModel(_Model):  # param.jl:237-311
    pass


_solver = None
_resident = {"token": None}


def _get_solver(device=0):
    global _solver
    if _solver is None or _solver.device != device:
        _solver = Solver(device)
    return _solver


def dc_EGM(p, model=_abi.MODEL_F, demo=False, device=0, strict=True):
    """dc_EGM!(m, p) of the reference for one Param; returns the flat Solution (dc_algo.jl:8-206)."""
    sol = _get_solver(device).solve(model, [p], keep_dirty=demo)
    _resident["token"] = sol  # the solution now resident on the device (sim() reuses it)
    if strict:
        _raise_for_status(int(sol.status[0]))
    return sol


def runf(par=None, device=0, strict=True, **kw):
    p = Param(par=par, **kw)
    return FModel(p, dc_EGM(p, _abi.MODEL_F, False, device, strict), _abi.MODEL_F), p


def runfdemo(par=None, device=0, strict=True, **kw):
    p = Param(par=par, **kw)
    return FModel(p, dc_EGM(p, _abi.MODEL_F, True, device, strict), _abi.MODEL_F), p


def rung(par=None, device=0, strict=True, **kw):
    p = Param(par=par, **kw)
    return GModel(p, dc_EGM(p, _abi.MODEL_G, False, device, strict), _abi.MODEL_G), p


def minimal_EGM(p=None, device=0):
    """minimal_EGM(p) (minimal_EGM.jl:3-43).  The reference returns only a plot; like
    minimal_EGM_borrow (minimal_EGM.jl:88) this returns the data: lists m[it], c[it], it = 1..nT."""
    p = p or Param()
    sol = _get_solver(device).solve(_abi.MODEL_MINIMAL, [p])
    _raise_for_status(int(sol.status[0]))
    m, c = [], []
    for it in range(p.nT):
        x, y = sol.line(0, _abi.KIND_C, 0, it)
        m.append(x)
        c.append(y)
    return m, c


def _minimal_lines(model, p, device, kinds):
    p = p or Param()
    sol = _get_solver(device).solve(model, [p])
    _resident["token"] = sol
    _raise_for_status(int(sol.status[0]))
    out = [[] for _ in range(1 + len(kinds))]
    for it in range(p.nT):
        x, y = sol.line(0, kinds[0], 0, it)
        out[0].append(x)
        out[1].append(y)
        for j, k in enumerate(kinds[1:]):
            out[2 + j].append(sol.line(0, k, 0, it)[1])
    return tuple(out)


def minimal_EGM_borrow(p=None, device=0):
    """minimal_EGM_borrow(p) (minimal_EGM.jl:49-89): returns (m, c) — the reference's third return value is the plot."""
    return _minimal_lines(_abi.MODEL_MINIMAL_BORROW, p, device, (_abi.KIND_C,))


def minimal_EGM_bequest(p=None, device=0):
    """minimal_EGM_bequest(p) (minimal_EGM.jl:92-159).  The reference returns only the two-panel plot; this
    returns the plotted data (m, c, v), lists over it = 1..nT."""
    return _minimal_lines(_abi.MODEL_MINIMAL_BEQUEST, p, device, (_abi.KIND_C, _abi.KIND_V))


def bm(samples=39, device=0, verbose=True):
    """bm() (bench.jl:3-31) without the MATLAB shell-out: one warm-up runf, then `samples` timed
    runf(par=pd) calls end to end (host setup, H2D, kernel, D2H of every line).  Returns the timing
    summary BenchmarkTools would print (the reference discards it, bench.jl:29)."""
    pd = dict(na=500, beta=0.95, sigma=0.35, R=1.05)
    pd["lambda"] = 0.000002
    runf(par=pd, device=device)
    ts = []
    for _ in range(samples):
        t0 = time.perf_counter()
        runf(par=pd, device=device)
        ts.append(1e3 * (time.perf_counter() - t0))
    ts = np.array(ts)
    out = dict(samples=samples, min_ms=float(ts.min()), median_ms=float(np.median(ts)), mean_ms=float(ts.mean()),
               max_ms=float(ts.max()), kernel_ms=float(_get_solver(device).timing().kernel_ms))
    if verbose:
        print("dcegm-b200 timing (runf, na=500, 25 periods):")
        print("  minimum time: %(min_ms).3f ms\n  median time:  %(median_ms).3f ms\n  mean time:    %(mean_ms).3f ms\n"
              "  maximum time: %(max_ms).3f ms\n  samples: %(samples)d  (kernel only: %(kernel_ms).3f ms)" % out)
    return out


class Simulation:  # simulate.jl:1-28
    """Panel of one simulated model: matrices (nsims, nT) w0, w1, cons, util, inc, prob_work (float),
    ystate, worker (int), ret_age (nsims), plus the per-period means the sweep summaries use."""

    def __init__(self, p, panel, moments, ret_hist):
        self.p = p
        for k, v in panel.items():
            setattr(self, k, v)
        self.moments = {k: moments[:, j] for j, k in enumerate(_abi.SIM_MOMENTS)}
        self.ret_hist = ret_hist


def sim(m, p, shocks=None, device=0, strict=True):
    """sim(m::FModel,p) / sim(m::GModel,p) (simulate.jl:61-203) on the GPU.  shocks=None draws the
    counter-based shocks keyed by p.rseed; pass the reference's raw rand() matrices to follow a Julia run."""
    model = _abi.MODEL_G if isinstance(m, GModel) else _abi.MODEL_F
    s = _get_solver(device)
    if _resident["token"] is not m.solution:  # another solve replaced it on the device
        m.solution = s.solve(model, [p])
        _resident["token"] = m.solution
    s.simulate(panel=True, shocks=shocks)
    mom, hist, st = s.sim_moments()
    if strict:
        _raise_for_status(int(st[0]))
    nsims = p.nsims if shocks is None else len(shocks["w0shocks"])
    return Simulation(p, s.sim_panel(0, nsims), mom[0], hist[0])


def rs(par=None, **kw):  # simulate.jl:271-274
    m, p = runf(par=par, **kw)
    return sim(m, p)


def rsg(par=None, **kw):  # simulate.jl:279-282
    m, p = rung(par=par, **kw)
    return sim(m, p)


def _u(x, working, p):  # utils.jl:40-46 (host-side, plotting helper only)
    x = np.asarray(x, float)
    with np.errstate(divide="ignore"):  # u(0) = -Inf, as in the reference
        if p.gamma == 1.0:
            return np.log(x) - p.alpha * working
        return (1.0 / (1.0 - p.gamma)) * (x ** (1.0 - p.gamma) - 1.0) - p.alpha * working


def v_analytic(m, p, id, it, iy=None):
    """v_analytic(m,p,id,it[,iy]) (plotting.jl:3-42): the value function of choice `id` with the
    credit-constrained region filled in analytically (u(c) + beta*vbound on a log-spaced grid of p.k points
    below the second knot), joined to the stored envelope at its second point.  Returns an (n, 2) array of
    (cash-on-hand, value).  Host-side post-processing of a fetched solution, like the reference's plotting code."""
    from .io import scale_grid

    g = isinstance(m, GModel)
    vf = m.v[(id, iy, it)] if g else m.v[(id, it)]
    c = m.c[(id, iy, it)] if g else m.c[(id, it)]
    vx, vy = getx(vf), gety(vf)
    if it == p.nT and not g:  # plotting.jl:6-9
        cash = np.linspace(p.a_low, p.a_high, p.na)
        if p.nu > 0:
            vals = p.nu * _u_plain(cash + p.bbar, p)
        else:
            vals = _u(cash, id == 1, p) + p.beta * bound(vf)
        return np.column_stack([cash, vals])
    cons = scale_grid(p.cfloor_plot, gety(c)[1], p.k, 1)   # plotting.jl:10, 34
    cash = scale_grid(p.a_low, vx[1], p.k, 1)              # plotting.jl:11, 35
    if it == p.nT and p.nu > 0:
        vals = p.nu * _u_plain(cons + p.bbar, p)
    else:
        mask = (cash < vx[1]) | (it == p.nT)               # vfun, dc_algo.jl:238-250
        vals = np.where(mask, _u(cons, id == 1, p) + p.beta * bound(vf), np.interp(cash, vx, vy))
    return np.vstack([np.column_stack([cash, vals]), np.column_stack([vx[1:], vy[1:]])])


def _u_plain(x, p):  # utils.jl:23-29
    x = np.asarray(x, float)
    return np.log(x) if p.gamma == 1.0 else (1.0 / (1.0 - p.gamma)) * x ** (1.0 - p.gamma)
