"""End-to-end run of the fine-grid configuration through dcegm_solve_f_packed: host Param array in, every stored
polyline densely packed in pinned host memory out (chunked solve overlapped with packing and D2H).
usage: c4_e2e.py [nmodels]"""
import ctypes as C
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

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
rng = np.random.default_rng(25032020)
base = Param(par=dict(na=20000, nT=50, ny=32, retage=51, beta=0.95, R=1.05)).to_c()
arr = params_array([base] * n)
lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), n)
sig = rng.uniform(0, 0.6, n)
for i in range(n):
    arr[i].lambda_, arr[i].sigma = float(lam[i]), float(sig[i])
with Solver(0) as s:
    lib, h = s.lib, s.h
    s.upload(_abi.MODEL_F, arr)
    s.run()
    kernel_ms = s.timing().kernel_ms
    lay = s.layout()
    nslots = n * lay.nT * lay.nstate * 4
    offsets = np.zeros(nslots + 1, np.int64)
    op = offsets.ctypes.data_as(C.POINTER(C.c_int64))
    npts = C.c_int64()
    lib.dcegm_fetch_packed(h, 0, n, None, 0, op, C.byref(npts))
    cap_pts = int(npts.value * 1.02) + 1024
    pinned = C.c_void_p()
    assert lib.dcegm_host_alloc(C.byref(pinned), cap_pts * 16) == 0
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        rc = lib.dcegm_solve_f_packed(h, arr, n, 0, C.cast(pinned, C.POINTER(C.c_double)), cap_pts, op, C.byref(npts))
        assert rc == 0, lib.dcegm_last_error(h)
        ts.append(time.perf_counter() - t0)
    # the same in the compact line format (dcegm_solve_f_compact): ordinates + abscissa runs
    woff = np.zeros(nslots + 1, np.int64)
    lens = np.zeros(nslots, np.int32)
    aux = np.zeros(nslots, np.int32)
    nw = C.c_int64()
    lp, ip = C.POINTER(C.c_int64), C.POINTER(C.c_int32)
    tc = []
    for _ in range(3):
        t0 = time.perf_counter()
        rc = lib.dcegm_solve_f_compact(h, arr, n, 0, C.cast(pinned, C.POINTER(C.c_double)), cap_pts * 2, woff.ctypes.data_as(lp),
                                       lens.ctypes.data_as(ip), aux.ctypes.data_as(ip), C.byref(nw))
        assert rc == 0, lib.dcegm_last_error(h)
        tc.append(time.perf_counter() - t0)
    st = s.status()
    ok = int(((st.astype(np.int64) & _abi.ST_FATAL_MASK) == 0).sum())
    lib.dcegm_host_free(pinned)
    print(json.dumps({"config": f"C4 e2e compact: FModel x{n} through dcegm_solve_f_compact", "e2e_ms": min(tc[1:]) * 1e3,
                      "d2h_bytes": int(nw.value * 8), "solved": ok, "models_per_s_e2e_solved": ok / min(tc[1:])}))
    print(json.dumps({"config": f"C4 e2e: FModel x{n} (na=20000, nT=50, ny=32) through dcegm_solve_f_packed",
                      "resident_solve_ms": kernel_ms, "e2e_ms": min(ts[1:]) * 1e3, "d2h_bytes": int(npts.value * 16),
                      "d2h_gbs": npts.value * 16 / min(ts[1:]) / 1e9, "models_per_s_e2e": n / min(ts[1:])}))
