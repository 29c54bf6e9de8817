"""Workload for ncu captures of the compact encoder alone: solve a C3 batch, then dcegm_fetch_compact on the resident
solution (compact_count_kernel -> host prefix -> compact_write_kernel).  usage: compact_only.py [nmodels]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench_configs import c3_params  # noqa: E402
from dcegm_jl_b200 import _abi  # noqa: E402
from dcegm_jl_b200.solver import Solver  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 592
with Solver(0) as s:
    s.solve(_abi.MODEL_F, c3_params(n), fetch=False)
    for _ in range(2):
        cl = s.fetch_compact()
    print("models", n, "compact MB", cl.nbytes / 1e6)
