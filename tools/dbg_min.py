import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from oracle import oracle
from dcegm_jl_b200 import Param, _abi
from dcegm_jl_b200.solver import Solver
s = Solver(0)
ps = [Param(par=dict(na=500, ny=9, nT=40, sigma=0.2)), Param(par=dict(na=500, ny=9, nT=40, beta=0.99))]
g = s.solve(_abi.MODEL_MINIMAL, ps); r = oracle.solve_minimal(ps)
print("lens eq", np.array_equal(g.lens, r.lens), g.status, r.status)
bad = np.argwhere(g.lens != r.lens); print(bad[:5])
d = g.lines[:, :, 0, 0, 0, :501] != r.lines[:, :, 0, 0, 0, :501]
print("ndiff", d.sum(), np.argwhere(d)[:5])
if d.any():
    i = tuple(np.argwhere(d)[0]); print(g.lines[:, :, 0, 0, 0, :501][i], r.lines[:, :, 0, 0, 0, :501][i])
