"""Census of status-flagged models of the BASELINE configurations: which bit, in which period, on which line, the
first offending abscissae and their cause.  Works on the CPU oracle (tests' checker; C3/C5 sizes) or on the GPU
library (any size; C4 needs it).  One JSON summary line per configuration + a table of examples on stderr.

usage: flag_census.py {c3|c4|c5} [--backend oracle|gpu] [--n N] [--chunk K] [--threads T] [--examples E]
"""
import argparse
import json
import os
import sys
from collections import Counter

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dcegm_jl_b200 import Param, _abi  # noqa: E402
from dcegm_jl_b200.param import params_array  # noqa: E402

BITS = {b: n for n, b in vars(_abi).items() if n.startswith("ST_") and isinstance(b, int) and b and (b & (b - 1)) == 0}


def draw(config, n):
    """the synthetic draws of SURVEY.md §8(d) (same generators as bench.py / tools/config_runs.py)"""
    rng = np.random.default_rng(25032020)
    if config == "c3":
        lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), size=n)
        sig = rng.uniform(0, 0.6, size=n)
        base = Param(par=dict(na=500, beta=0.95, R=1.05, nT=25, ny=5)).to_c()
        arr = params_array([base] * n)
        for i in range(n):
            arr[i].lambda_, arr[i].sigma = float(lam[i]), float(sig[i])
        return _abi.MODEL_F, arr
    if config == "c4":
        base = Param(par=dict(na=20000, nT=50, ny=32, retage=51, beta=0.95, R=1.05)).to_c()
        arr = params_array([base] * n)
        lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), n)
        sig = rng.uniform(0, 0.6, n)
        for i in range(n):
            arr[i].lambda_, arr[i].sigma = float(lam[i]), float(sig[i])
        return _abi.MODEL_F, arr
    if config == "c5":
        ps = []
        for _ in range(n):
            d = dict(na=100, nsims=1000, beta=float(rng.uniform(0.5, 1)), R=float(rng.uniform(1, 1.5)),
                     alpha=float(rng.uniform(0, 0.5)), sigma=float(rng.uniform(0, 0.35)), rho=float(rng.uniform(0, 0.95)),
                     pension=float(rng.uniform(0, 1)))
            d["lambda"] = float(10 ** rng.uniform(np.log10(2e-7), 0))
            ps.append(Param(par=d))
        return _abi.MODEL_G, params_array(ps)
    raise SystemExit("config must be c3, c4 or c5")


LINE_NAMES = {(0, 0): "worker policy", (0, 1): "worker value", (1, 0): "retiree policy", (1, 1): "retiree value"}


def first_defect(sol, m):
    """scan the periods backwards (the order of the solve): the first line that is not strictly increasing"""
    lay = sol.lay
    for it in range(lay.nT - 2, -1, -1):
        for iy in range(lay.nstate):
            for (id_, kind), name in LINE_NAMES.items():
                n = int(sol.lens[m, it, iy, id_, kind])
                x = sol.lines[m, it, iy, id_, kind, :n, 0]
                d = np.diff(x)
                bad = np.nonzero(~(d > 0))[0]
                if bad.size == 0:
                    continue
                k = int(bad[0])
                ni = int(sol.meta[m, it, iy, 1])
                isx = sol.isects[m, it, iy, :ni, 0]
                xa, xb = float(x[k]), float(x[k + 1])
                if xa == xb:
                    cause = "duplicate abscissa"
                    if id_ == 0 and ni and (np.any(isx == xa)):
                        cause += ": intersection at a turning point shared by two sections (envelope.jl:199 f(pt1)==0; :212 and :248-252 push the point twice)"
                    elif id_ == 0 and kind == 0 and ni and np.any(np.abs(isx - xa) <= 3e-13):
                        cause += ": kink pair (x*-1e3 eps, x*) meets a surviving point (dc_algo.jl:486-492)"
                elif not np.isfinite(xa) or not np.isfinite(xb):
                    cause = "non-finite abscissa"
                else:
                    cause = "decreasing abscissae"
                    if id_ == 1:
                        cause += ": the retiree's endogenous grid a + c0 bends back (no envelope is run for id = 2, dc_algo.jl:87-92)"
                    elif kind == 0 and ni and np.min(np.abs(isx - xb)) <= 3e-13 * max(1.0, abs(xb)):
                        cause += ": inserted kink pair out of order"
                return dict(it=it + 1, iy=iy + 1, line=name, k=k, x_k=xa, x_k1=xb, n=n, n_isects=ni, n_bad=int(bad.size), cause=cause)
    return None


def census(sol, params, first, examples, out):
    for m in range(sol.lay.nmodels):
        st = int(sol.status[m]) & 0xFFFFFFFF
        out["models"] += 1
        if st == 0:
            continue
        for b, name in BITS.items():
            if st & b:
                out["bits"][name] += 1
        if st & _abi.ST_FATAL_MASK:
            out["fatal"] += 1
        else:
            out["info_only"] += 1
        rec = dict(model=first + m, status=hex(st), sigma=params[first + m].sigma, lam=params[first + m].lambda_)
        if st & (_abi.ST_NONMONOTONE | _abi.ST_NONMONO_FIRST | _abi.ST_DUP_KNOT):
            d = first_defect(sol, m)
            if d:
                rec.update(d)
                out["causes"][d["cause"]] += 1
                out["lines"][d["line"]] += 1
                out["periods"][d["it"]] += 1
        if st & (_abi.ST_MISALIGNED | _abi.ST_EXTRA_RIGHT | _abi.ST_INSERT_RANGE):
            # periods whose worker envelope produced intersections but a different number of inserted pairs
            lay = sol.lay
            for it in range(lay.nT - 2, -1, -1):
                for iy in range(lay.nstate):
                    ni, nrem = int(sol.meta[m, it, iy, 1]), int(sol.meta[m, it, iy, 2])
                    nc = int(sol.lens[m, it, iy, 0, 0])
                    npre = int(sol.meta[m, it, iy, 3])
                    if nrem and nc - 1 != lay.na + npre - nrem + 2 * ni:
                        rec.setdefault("pair_mismatch", dict(it=it + 1, iy=iy + 1, n_isects=ni, n_removed=nrem, len_c=nc,
                                                             expected=lay.na + npre - nrem + 2 * ni + 1))
        if len(out["examples"]) < examples:
            out["examples"].append(rec)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config")
    ap.add_argument("--backend", default="oracle")
    ap.add_argument("--n", type=int, default=0)
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--examples", type=int, default=12)
    ap.add_argument("--portable", type=int, default=1, help="oracle: dc_math.h arithmetic (bit-equal to the GPU) or libm")
    a = ap.parse_args()
    n = a.n or {"c3": 4096, "c4": 256, "c5": 8192}[a.config]
    chunk = a.chunk or {"c3": 512, "c4": 8, "c5": 1024}[a.config]
    model, arr = draw(a.config, n)
    out = dict(config=a.config, backend=a.backend, models=0, fatal=0, info_only=0, bits=Counter(), causes=Counter(),
               lines=Counter(), periods=Counter(), examples=[])
    solver = None
    if a.backend == "gpu":
        from dcegm_jl_b200.solver import Solver

        solver = Solver(0)
        solver.solve(model, arr, fetch=False)
        for c0 in range(0, n, chunk):
            cnt = min(chunk, n - c0)
            census(solver.fetch(c0, cnt), arr, c0, a.examples, out)
        solver.close()
    else:
        from oracle import oracle

        for c0 in range(0, n, chunk):
            cnt = min(chunk, n - c0)
            sub = params_array([arr[i] for i in range(c0, c0 + cnt)])
            census(oracle.solve(model, sub, nthreads=a.threads, portable=bool(a.portable)), arr, c0, a.examples, out)
    ex = out.pop("examples")
    for k in ("bits", "causes", "lines", "periods"):
        out[k] = dict(sorted(out[k].items(), key=lambda kv: str(kv[0])))
    print(json.dumps(out))
    for r in ex:
        print(json.dumps(r), file=sys.stderr)


if __name__ == "__main__":
    main()
