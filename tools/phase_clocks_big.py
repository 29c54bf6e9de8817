"""Per-phase cycle shares of big_env_kernel (fine-grid solver) from the profiling build
(make libdcegm_cuda_prof.so in dcegm.jl_b200/csrc).  usage: phase_clocks_big.py [nmodels] [na] [ny] [nT]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dcegm_jl_b200 import _lib  # noqa: E402

_lib.LIB_PATH = os.path.join(ROOT, "dcegm.jl_b200", "csrc", "libdcegm_cuda_prof.so")
from dcegm_jl_b200 import Param, _abi  # noqa: E402
from dcegm_jl_b200.param import params_array  # noqa: E402
from dcegm_jl_b200.solver import Solver  # noqa: E402

a = [int(x) for x in sys.argv[1:]]
n, na, ny, nT = (a + [256, 20000, 32, 50][len(a):])[:4]
NAMES = {4: "env: bend-back guard", 5: "env: split + removed->global + repair prep", 6: "env: every point",
         7: "env: sort/unique", 8: "env: switches", 9: "env: emission/removed", 10: "env: policy repair",
         12: "env: tail checks", 13: "retiree tail + store", 14: "knot index / summary", 0: "dirty store"}
rng = np.random.default_rng(25032020)
base = Param(par=dict(na=na, nT=nT, ny=ny, retage=nT + 1, beta=0.95, R=1.05)).to_c()
arr = params_array([base] * n)
lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), n)
sig = rng.uniform(0, 0.6, n)
for i in range(n):
    arr[i].lambda_, arr[i].sigma = float(lam[i]), float(sig[i])
s = Solver(0)
s.upload(_abi.MODEL_F, arr, fine_grid=True)
s.run()
out = (C.c_ulonglong * 16)()
s.lib.dcegm_debug_phase_clocks_big(out, 1)
s.run()
s.lib.dcegm_debug_phase_clocks_big(out, 1)
tot = sum(out)
print(f"solve {s.timing().kernel_ms:.3f} ms for {n} models na={na} ny={ny} nT={nT} (instrumented build), "
      f"flagged {int((s.status() != 0).sum())}")
for i, v in enumerate(out):
    if v:
        print(f"{i:2d} {NAMES.get(i, '-'):44s} {100.0 * v / tot:6.2f}%")
