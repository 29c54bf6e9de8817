"""On-disk format of the reference's golden data and the F_test comparison on top of it.

matlab/polyline.m:472-484 (`write`) stores one polyline per file, `<name>_<id>_<it>.csv`, two rows
(x; y), comma separated, 8 significant digits; matlab/bench.m:13-26 writes `policy` and `value` for
id = 1,2 and it = 1..T, and test/F_test.jl:47-69 reads them back with readdlm and compares them with the
Julia solution.  This module writes a solved FModel in that format, reads such a directory, and runs
F_test's checks — so a solution from this library can be diffed against MATLAB/Julia output files
wherever those exist.  Pure host-side post-processing of fetched lines.
"""
import os

import numpy as np

from . import _abi

KINDS = {"policy": _abi.KIND_C, "value": _abi.KIND_V}


def _fmt(v):
    # dlmwrite(..., 'precision', 8) prints with %.8g
    return "%.8g" % v


def write_csv(m, outdir, names=("policy", "value")):
    """write m.c / m.v of a solved FModel as <outdir>/{policy,value}_<id>_<it>.csv (1-based id, it)"""
    os.makedirs(outdir, exist_ok=True)
    sol = m.solution
    nT = sol.lay.nT
    files = []
    for name in names:
        for id_ in (1, 2):
            for it in range(1, nT + 1):
                x, y = sol.line(0, KINDS[name], id_ - 1, it - 1)
                path = os.path.join(outdir, f"{name}_{id_}_{it}.csv")
                with open(path, "w") as f:
                    f.write(",".join(_fmt(v) for v in x) + "\n")
                    f.write(",".join(_fmt(v) for v in y) + "\n")
                files.append(path)
    return files


def read_csv(outdir, nT, names=("policy", "value")):
    """{(name, id, it): array (2, n)} like readdlm(..., ',') in F_test.jl:49-50"""
    out = {}
    for name in names:
        for id_ in (1, 2):
            for it in range(1, nT + 1):
                a = np.loadtxt(os.path.join(outdir, f"{name}_{id_}_{it}.csv"), delimiter=",", ndmin=2)
                out[(name, id_, it)] = a
    return out


def f_test(m, golden, tol=1e-4, tol_vy=5e-4):
    """test/F_test.jl:47-69 on a solved FModel `m` against `golden` (dict from read_csv, or an npz with
    keys '<name>_<id>_<it>').  Returns the list of failed checks (empty = the reference's test passes)."""
    sol = m.solution
    nT = sol.lay.nT
    get = (lambda n, i, t: golden[(n, i, t)]) if not hasattr(golden, "files") else (lambda n, i, t: golden[f"{n}_{i}_{t}"])
    bad = []
    for i in (1, 2):
        for it in range(nT, 0, -1):
            mpo, mvf = get("policy", i, it), get("value", i, it)
            cx, cy = sol.line(0, _abi.KIND_C, i - 1, it - 1)
            vx, vy = sol.line(0, _abi.KIND_V, i - 1, it - 1)
            if len(cx) != mpo.shape[1] or len(vx) != mvf.shape[1]:  # F_test.jl:53-54
                bad.append((i, it, "length", len(cx), mpo.shape[1], len(vx), mvf.shape[1]))
                continue
            srt = np.argsort(mpo[0], kind="stable")  # F_test.jl:55
            if np.max(np.abs(cx - mpo[0, srt])) >= tol:
                bad.append((i, it, "policy x"))
            if np.max(np.abs(cy - mpo[1, srt])) >= tol:
                bad.append((i, it, "policy y"))
            if it == nT:  # F_test.jl:59-61
                if abs(vx[0] - mvf[0, 0]) > tol or abs(vy[0] - mvf[1, 0]) > tol:
                    bad.append((i, it, "terminal value"))
            else:
                if np.max(np.abs(vx - mvf[0])) >= tol:
                    bad.append((i, it, "value x"))
                if np.max(np.abs(vy - mvf[1])) >= tol_vy:
                    bad.append((i, it, "value y"))
    return bad


def scale_grid(lb, ub, n, logorder=1, offset=0.05):
    """scaleGrid (utils.jl:200-238), used by v_analytic for its plotting grids"""
    if lb >= ub:
        lb = ub - abs(ub) * offset
    if logorder == 0:
        return np.linspace(lb, ub, n)
    off = 1.0 - lb if lb < 0 else 1.0
    f, g = lb, ub
    for _ in range(logorder):
        f, g = np.log(f + off), np.log(g + off)
    out = np.linspace(f, g, n)
    for _ in range(logorder):
        out = np.exp(out) - off
    return out
