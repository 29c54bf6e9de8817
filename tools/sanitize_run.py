"""Small workload for compute-sanitizer (memcheck / racecheck / synccheck): a few C3 retirement models on the
resident solver (small lambda: many sections, intersections, bend-back), the same models on the fine-grid solver,
general models + simulation, minimal EGM.  Run as
    compute-sanitizer --tool racecheck python tools/sanitize_run.py [part ...]
parts: f (resident F), big (forced fine grid), g (general model + sim), min (minimal EGM); default all."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dcegm_jl_b200 import Param, _abi  # noqa: E402
from dcegm_jl_b200.solver import Solver  # noqa: E402

parts = sys.argv[1:] or ["f", "big", "g", "min"]


def fparams(na):
    ps = []
    for lam, sig in ((2e-6, 0.35), (1e-4, 0.0), (1e-3, 0.1), (2e-6, 0.045), (0.05, 0.5), (1e-2, 0.6)):
        d = dict(na=na, beta=0.95, R=1.05, sigma=sig)
        d["lambda"] = lam
        ps.append(Param(par=d))
    return ps


with Solver(0) as s:
    if "f" in parts:
        sol = s.solve(_abi.MODEL_F, fparams(500))
        print("resident F: status", [hex(int(x) & 0xFFFFFFFF) for x in sol.status], "isects", int(sol.meta[..., 1].sum()))
    if "big" in parts:
        sol = s.solve(_abi.MODEL_F, fparams(500)[:3], fine_grid=True)
        print("fine-grid F: status", [hex(int(x) & 0xFFFFFFFF) for x in sol.status], "isects", int(sol.meta[..., 1].sum()))
    if "g" in parts:
        pg = [Param(par=dict(na=100, nsims=2000)), Param(par=dict(na=100, nsims=2000, rho=0.3, sigma=0.2, alpha=0.1))]
        sol = s.solve(_abi.MODEL_G, pg)
        s.simulate()
        mom, hist, st = s.sim_moments()
        print("general model: status", [hex(int(x) & 0xFFFFFFFF) for x in sol.status], "agents", int(hist.sum()))
    if "min" in parts:
        sol = s.solve(_abi.MODEL_MINIMAL, [Param(), Param(par=dict(beta=0.9))])
        print("minimal: lens", sol.lens.reshape(-1)[:3])
print("sanitize_run done")
