"""C5-shard style run for profiling the forward simulator: n GModels (na=100, ny=5, nT=25) solved, then simulated
with nsims agents each.  usage: sim_run.py [nmodels] [nsims] [f|g]"""
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

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1184
nsims = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
kind = sys.argv[3] if len(sys.argv) > 3 else "g"
rng = np.random.default_rng(25032020)
ps = []
for _ in range(n):
    d = dict(na=100, nsims=nsims, beta=float(rng.uniform(0.5, 1)), R=float(rng.uniform(1, 1.5)),
             alpha=float(rng.uniform(0, 0.5)), sigma=float(rng.uniform(0, 0.35)), rho=float(rng.uniform(0, 0.95)),
             pension=float(rng.uniform(0, 1)))
    d["lambda"] = float(10 ** rng.uniform(np.log10(2e-7), 0))
    ps.append(Param(par=d))
arr = params_array(ps)
with Solver(0) as s:
    s.upload(_abi.MODEL_G if kind == "g" else _abi.MODEL_F, arr)
    s.run()
    s.simulate()
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        s.simulate()
        ts.append(time.perf_counter() - t0)
    mom, hist, st = s.sim_moments()
    print(json.dumps({"models": n, "nsims": nsims, "kind": kind, "solve_ms": s.timing().kernel_ms, "sim_s": min(ts),
                      "agent_periods_per_s": n * nsims * 25 / min(ts), "flagged_sim": int((st != 0).sum()),
                      "mean_cons_t1": float(np.nanmean(mom[:, 0, 2]))}))
