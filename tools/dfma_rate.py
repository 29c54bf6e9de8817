"""FP64 issue rate of cuda:0 (dcegm_debug_dfma_rate): thread-level DFMA/s of independent chains at full occupancy."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dcegm_jl_b200 import _lib  # noqa: E402

lib = _lib.load()
r, ms = C.c_double(), C.c_double()
rc = lib.dcegm_debug_dfma_rate(0, C.byref(r), C.byref(ms))
print(json.dumps({"rc": rc, "dfma_per_s": r.value, "tflops_fp64": 2 * r.value / 1e12, "warp_inst_per_s": r.value / 32, "ms": ms.value}))
