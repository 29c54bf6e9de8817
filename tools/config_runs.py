"""Times the other BASELINE.json configurations on one GPU (they are parity-test cases, not bench lines):
C2 minimal EGM x4096, C5 shard (GModel x8192, na=100) + forward simulation, optionally C4 (fine grid).
usage: config_runs.py [c2] [c5 [nsims]] [c4 [nmodels]]   -> one JSON line per configuration"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dcegm_jl_b200 import Param, _abi  # noqa: E402
from dcegm_jl_b200.param import params_array  # noqa: E402
from dcegm_jl_b200.solver import Solver  # noqa: E402

args = sys.argv[1:] or ["c2", "c5"]
rng = np.random.default_rng(25032020)


def solved(st):
    """models whose status carries no fatal bit: only these count as solved (informational bits do not matter)"""
    return int(((st.astype(np.int64) & _abi.ST_FATAL_MASK) == 0).sum())


def bits(st):
    names = {b: n for n, b in vars(_abi).items() if n.startswith("ST_") and isinstance(b, int) and b and (b & (b - 1)) == 0}
    return {names[1 << b]: int(((st >> b) & 1).sum()) for b in range(16) if ((st >> b) & 1).any()}


def timed(f, reps=3):
    f()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        f()
        ts.append(time.perf_counter() - t0)
    return min(ts)


with Solver(0) as s:
    if "c2" in args:
        n = 4096
        base = Param(par=dict(sigma=0.35, na=100, ny=5, nT=25, a_high=50.0, cfloor=1e-3)).to_c()
        arr = params_array([base] * n)
        for i, (b, r) in enumerate(zip(rng.uniform(0.5, 1.0, n), rng.uniform(1.0, 1.5, n))):
            arr[i].beta, arr[i].R = float(b), float(r)
        s.upload(_abi.MODEL_MINIMAL, arr)
        dt = timed(s.run)
        ok = solved(s.status())
        print(json.dumps({"config": "C2 minimal EGM x4096 (na=100, nT=25)", "kernel_ms": s.timing().kernel_ms,
                          "models_per_s": ok / (s.timing().kernel_ms * 1e-3), "wall_ms": dt * 1e3, "models": n,
                          "solved": ok, "status_bits": bits(s.status())}))
    if "c5" in args:
        k = args.index("c5")
        nsims = int(args[k + 1]) if len(args) > k + 1 and args[k + 1].isdigit() else 100000
        n = 8192
        ps = []
        for _ in range(n):
            d = dict(na=100, nsims=nsims, beta=float(rng.uniform(0.5, 1)), R=float(rng.uniform(1, 1.5)),
                     alpha=float(rng.uniform(0, 0.5)), sigma=float(rng.uniform(0, 0.35)), rho=float(rng.uniform(0, 0.95)),
                     pension=float(rng.uniform(0, 1)))
            d["lambda"] = float(10 ** rng.uniform(np.log10(2e-7), 0))
            ps.append(Param(par=d))
        arr = params_array(ps)
        s.upload(_abi.MODEL_G, arr)
        dt = timed(s.run)
        kms = s.timing().kernel_ms
        t0 = time.perf_counter()
        s.simulate()
        ts = time.perf_counter() - t0
        mom, hist, st = s.sim_moments()
        ok = solved(s.status())
        print(json.dumps({"config": f"C5 shard: GModel x{n} (na=100, ny=5, nT=25) + sim of {nsims} agents each",
                          "solve_kernel_ms": kms, "models_per_s": ok / (kms * 1e-3), "sim_s": ts, "models": n, "solved": ok,
                          "status_bits": bits(s.status()), "agent_periods_per_s_all_models": n * nsims * 25 / ts,
                          "flagged_sim": int((st != 0).sum()), "mean_cons_t1": float(np.nanmean(mom[:, 0, 2]))}))
    if "c4" in args:
        k = args.index("c4")
        n = int(args[k + 1]) if len(args) > k + 1 and args[k + 1].isdigit() else 256
        base = Param(par=dict(na=20000, nT=50, ny=32, retage=51, beta=0.95, R=1.05)).to_c()
        arr = params_array([base] * n)
        lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), n)
        sig = rng.uniform(0, 0.6, n)
        for i in range(n):
            arr[i].lambda_, arr[i].sigma = float(lam[i]), float(sig[i])
        s.upload(_abi.MODEL_F, arr)
        t0 = time.perf_counter()
        s.run()
        dt = time.perf_counter() - t0
        st = s.status()
        print(json.dumps({"config": f"C4 fine grid: FModel x{n} (na=20000, nT=50, ny=32)", "kernel_ms": s.timing().kernel_ms,
                          "models_per_s": solved(st) / dt, "models": n, "solved": solved(st), "status_bits": bits(st),
                          "launches": s.timing().n_launches}))
