"""Debug helper (GPU box): where do GPU and oracle differ most on the C3 parity batch?"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from oracle import oracle
from dcegm_jl_b200 import _abi
from dcegm_jl_b200.solver import Solver
from test_gpu_parity_f import c3_batch
np.set_printoptions(precision=17, linewidth=220)
ps = c3_batch(96)
s = Solver(0)
g = s.solve(_abi.MODEL_F, ps, keep_dirty=True); r = oracle.solve_f(ps, nthreads=16, keep_dirty=True)
lay = g.lay
k = np.arange(lay.cap)
mask = np.broadcast_to((k[None, None, None, None, None, :] < g.lens[..., None])[..., None], g.lines.shape)
err = np.where(mask, np.abs(g.lines - r.lines) / (0.01 + np.abs(r.lines)), 0.0)
err = np.nan_to_num(err)
permodel = err.max(axis=(1, 2, 3, 4, 5, 6))
print("models with err>1e-11:", np.nonzero(permodel > 1e-11)[0], permodel[permodel > 1e-11])
for _ in range(3):
    idx = np.unravel_index(err.argmax(), err.shape)
    m, it, iy, id, kind, pt, xy = idx
    print("worst", idx, err[idx], "lambda", ps[m].lambda_, "sigma", ps[m].sigma, "meta", g.meta[m, it, iy])
    lo = max(0, pt - 2)
    print(" gpu", g.lines[m, it, iy, id, kind, lo:pt + 3].tolist())
    print(" ref", r.lines[m, it, iy, id, kind, lo:pt + 3].tolist())
    print(" isects gpu", g.isect_points(m, it).tolist()); print(" isects ref", r.isect_points(m, it).tolist())
    d = np.abs(g.dirty[m] - r.dirty[m]) / (0.01 + np.abs(r.dirty[m]))
    print(" dirty err this model: max", d[:24].max(), "per period", d[:24].max(axis=(1, 2, 3, 4, 5)))
    # earliest period (highest it) where error exceeds 1e-12
    pe = err[m].max(axis=(1, 2, 3, 4, 5))
    print(" line err per period", pe)
    err[m] = 0
