"""Per-phase cycle shares of solve_f_kernel from the profiling build (make libdcegm_cuda_prof.so in
dcegm.jl_b200/csrc): every phase boundary is a block barrier + clock64 on thread 0, summed over CTAs.
usage: phase_clocks.py [batch]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dcegm_jl_b200 import _lib  # noqa: E402

_lib.LIB_PATH = os.path.join(ROOT, "dcegm.jl_b200", "csrc", "libdcegm_cuda_prof.so")
from bench_configs import c3_params  # noqa: E402
from dcegm_jl_b200 import _abi  # noqa: E402
from dcegm_jl_b200.solver import Solver  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1184
NAMES = ["period setup", "knot index build", "EGM worker", "EGM retiree", "env: bend-back guard", "env: split + repair prep",
         "env: every point", "env: sort/unique", "env: switches", "env: emission/removed", "env: policy repair", "-",
         "env: tail checks", "retiree tail + store", "summary", "-"]
s = Solver(0)
s.upload(_abi.MODEL_F, c3_params(n))
s.run()
out = (C.c_ulonglong * 16)()
s.lib.dcegm_debug_phase_clocks(out, 1)
s.run()
s.lib.dcegm_debug_phase_clocks(out, 1)
tot = sum(out)
print(f"kernel {s.timing().kernel_ms:.3f} ms for {n} models (instrumented build)")
for i, v in enumerate(out):
    if v:
        print(f"{i:2d} {NAMES[i]:28s} {100.0 * v / tot:6.2f}%")
