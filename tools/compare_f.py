"""Debug helper (GPU box): solve a few FModels on the GPU and with the oracle, print differences."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from oracle import oracle
from dcegm_jl_b200 import Param, _abi
from dcegm_jl_b200.solver import Solver
from parity import line_errors
from conftest import ftest_param, bm_param

s = Solver(0)
ps = [ftest_param(), bm_param(), ftest_param(na=100), bm_param(na=200, sigma=0.05, **{"lambda": 8e-4})]
for p in ps:
    t = time.time(); g = s.solve(_abi.MODEL_F, [p], keep_dirty=True); tg = time.time() - t
    r = oracle.solve_f([p], keep_dirty=True)
    print("na", p.na, "gpu ms", s.timing().kernel_ms, "wall", tg, "status", g.status, r.status)
    print("  lens equal", np.array_equal(g.lens, r.lens), "meta equal", np.array_equal(g.meta, r.meta),
          "removed equal", np.array_equal(g.removed, r.removed))
    if not np.array_equal(g.lens, r.lens):
        bad = np.argwhere(g.lens != r.lens)
        print("  first lens mismatches (m,it,iy,id,kind):", bad[:8].tolist(), g.lens[tuple(bad[0])], r.lens[tuple(bad[0])])
    if not np.array_equal(g.meta, r.meta):
        bad = np.argwhere(g.meta != r.meta)
        print("  meta mismatches:", bad[:8].tolist())
        for b in bad[:4]:
            print("   gpu", g.meta[b[0], b[1], b[2]], "ref", r.meta[b[0], b[1], b[2]])
    d = np.abs(g.dirty[:, :-1] - r.dirty[:, :-1]) / (1e-2 + np.abs(r.dirty[:, :-1]))
    print("  dirty max rel err", d.max(), "at", np.unravel_index(d.argmax(), d.shape))
    if np.array_equal(g.lens, r.lens):
        print("  line max rel err", line_errors(g, r))
