"""Timeline of one end-to-end C3 call (dcegm_solve_f_compact, host Param array in -> compact lines in pinned host
memory): the library's DCEGM_TRACE host timestamps of the last of several calls, plus wall time of each call.
usage: e2e_trace.py [nmodels] [packed]     packed: dcegm_solve_f_packed (every point as an (x, y) pair) instead"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench_configs import c3_params  # noqa: E402
from dcegm_jl_b200 import _abi  # noqa: E402
from dcegm_jl_b200.solver import Solver  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
arr = c3_params(n)
with Solver(0) as s:
    lib, h = s.lib, s.h
    s.upload(_abi.MODEL_F, arr)
    s.run()
    print("resident kernel_ms", s.timing().kernel_ms)
    lay = s.layout()
    nslots = n * lay.nT * lay.nstate * 4
    woff = np.zeros(nslots + 1, np.int64)
    lens = np.zeros(nslots, np.int32)
    aux = np.zeros(nslots, np.int32)
    nw = C.c_int64()
    lp, ip = C.POINTER(C.c_int64), C.POINTER(C.c_int32)
    args = (woff.ctypes.data_as(lp), lens.ctypes.data_as(ip), aux.ctypes.data_as(ip), C.byref(nw))
    assert lib.dcegm_fetch_compact(h, 0, n, None, 0, *args) == 0
    cap_words = int(nw.value * 1.05) + 4096
    pinned = C.c_void_p()
    assert lib.dcegm_host_alloc(C.byref(pinned), cap_words * 8) == 0
    packed = len(sys.argv) > 2 and sys.argv[2] == "packed"
    if packed:
        npts_total = int(lens.astype(np.int64).sum())
        lib.dcegm_host_free(pinned)
        assert lib.dcegm_host_alloc(C.byref(pinned), (npts_total + 256) * 16) == 0
    for rep in range(6):
        if rep == 5:
            os.environ["DCEGM_TRACE"] = "1"
        t0 = time.perf_counter()
        if packed:
            rc = lib.dcegm_solve_f_packed(h, arr, n, 0, C.cast(pinned, C.POINTER(C.c_double)), npts_total + 256, args[0], C.byref(nw))
        else:
            rc = lib.dcegm_solve_f_compact(h, arr, n, 0, C.cast(pinned, C.POINTER(C.c_double)), cap_words, *args)
        dt = time.perf_counter() - t0
        assert rc == 0, lib.dcegm_last_error(h)
        t = s.timing()
        print(f"call {rep}: {dt * 1e3:.2f} ms wall, {n / dt:.0f} models/s, {nw.value * 8 / 1e6:.1f} MB, kernel_ms {t.kernel_ms:.2f}, "
              f"setup {t.setup_host_ms:.2f}, h2d {t.h2d_ms:.2f}", flush=True)
    lib.dcegm_host_free(pinned)
