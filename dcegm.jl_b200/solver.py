"""Solver — thin owner of one library handle (one per GPU), returning Solution views."""
import ctypes as C

import numpy as np

from . import _abi, _lib
from .param import params_array
from .solution import Solution


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class Solver:
    def __init__(self, device=0):
        self.lib = _lib.load()
        self.h = C.c_void_p()
        rc = self.lib.dcegm_create(int(device), C.byref(self.h))
        if rc == _abi.E_NODEVICE:
            raise _lib.DcegmError(rc, "no CUDA device; libdcegm_cuda has no CPU path")
        if rc != 0:
            raise _lib.DcegmError(rc, "dcegm_create failed")
        self.device = device

    def close(self):
        if self.h:
            self.lib.dcegm_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc != 0:
            raise _lib.DcegmError(rc, self.lib.dcegm_last_error(self.h).decode())

    # ---- solving -------------------------------------------------------------------------
    def solve(self, model, params, keep_dirty=False, summary_only=False, fetch=True, fine_grid=False):
        """fine_grid: force the fine-grid retirement solver (automatic when the lines of one model do not fit
        into shared memory, i.e. na > ~1000)."""
        arr = params if isinstance(params, C.Array) else params_array(params)
        flags = ((_abi.KEEP_DIRTY if keep_dirty else 0) | (_abi.SUMMARY_ONLY if summary_only else 0)
                 | (_abi.FORCE_FINE_GRID if fine_grid else 0))
        fn = {_abi.MODEL_F: self.lib.dcegm_solve_f, _abi.MODEL_G: self.lib.dcegm_solve_g,
              _abi.MODEL_MINIMAL: self.lib.dcegm_solve_minimal, _abi.MODEL_MINIMAL_BORROW: self.lib.dcegm_solve_minimal_borrow,
              _abi.MODEL_MINIMAL_BEQUEST: self.lib.dcegm_solve_minimal_bequest}[model]
        self._ck(fn(self.h, arr, len(arr), flags))
        return self.fetch() if fetch and not summary_only else None

    def upload(self, model, params, keep_dirty=False, summary_only=False, fine_grid=False):
        arr = params if isinstance(params, C.Array) else params_array(params)
        flags = ((_abi.KEEP_DIRTY if keep_dirty else 0) | (_abi.SUMMARY_ONLY if summary_only else 0)
                 | (_abi.FORCE_FINE_GRID if fine_grid else 0))
        fn = {_abi.MODEL_F: self.lib.dcegm_upload_f, _abi.MODEL_G: self.lib.dcegm_upload_g,
              _abi.MODEL_MINIMAL: self.lib.dcegm_upload_minimal}[model]
        self._ck(fn(self.h, arr, len(arr), flags))

    def run(self):
        self._ck(self.lib.dcegm_run(self.h))

    def layout(self):
        lay = _abi.dcegm_layout()
        self._ck(self.lib.dcegm_get_layout(self.h, C.byref(lay)))
        return lay

    def timing(self):
        t = _abi.dcegm_timing()
        self._ck(self.lib.dcegm_last_timing(self.h, C.byref(t)))
        return t

    # ---- results ----------------------------------------------------------------------------
    def fetch(self, first=0, count=None, into=None):
        lay = self.layout()
        count = lay.nmodels - first if count is None else count
        sub = _abi.dcegm_layout()
        C.memmove(C.byref(sub), C.byref(lay), C.sizeof(lay))
        sub.nmodels = count
        s = into if into is not None else Solution.allocate(sub, bool(lay.has_dirty))
        if lay.has_lines:
            self._ck(self.lib.dcegm_fetch_lines(self.h, first, count, _dp(s.lines)))
        self._ck(self.lib.dcegm_fetch_lens(self.h, first, count, s.lens.ctypes.data_as(C.POINTER(C.c_int32))))
        self._ck(self.lib.dcegm_fetch_vbound(self.h, first, count, _dp(s.vbound)))
        self._ck(self.lib.dcegm_fetch_meta(self.h, first, count, s.meta.ctypes.data_as(C.c_void_p)))
        self._ck(self.lib.dcegm_fetch_isects(self.h, first, count, _dp(s.isects)))
        self._ck(self.lib.dcegm_fetch_removed(self.h, first, count, s.removed.ctypes.data_as(C.POINTER(C.c_uint32))))
        if lay.has_dirty:
            self._ck(self.lib.dcegm_fetch_dirty(self.h, first, count, _dp(s.dirty)))
        st = np.zeros(lay.nmodels, np.int32)
        self._ck(self.lib.dcegm_status(self.h, st.ctypes.data_as(C.POINTER(C.c_int32))))
        s.status[:] = st[first:first + count]
        return s

    def fetch_line(self, model, kind, id, it, iy=0):
        lay = self.layout()
        buf = np.zeros((max(lay.cap, lay.na), 2))
        n = C.c_int()
        self._ck(self.lib.dcegm_fetch_line(self.h, model, kind, id, iy, it, _dp(buf), buf.shape[0], C.byref(n)))
        return buf[: n.value, 0].copy(), buf[: n.value, 1].copy()

    def fetch_packed(self, first=0, count=None):
        lay = self.layout()
        count = lay.nmodels - first if count is None else count
        nsl = count * lay.nT * lay.nstate * 4
        offsets = np.zeros(nsl + 1, np.int64)
        npts = C.c_int64()
        op = offsets.ctypes.data_as(C.POINTER(C.c_int64))
        self._ck(self.lib.dcegm_fetch_packed(self.h, first, count, None, 0, op, C.byref(npts)))
        xy = np.zeros((npts.value, 2))
        self._ck(self.lib.dcegm_fetch_packed(self.h, first, count, _dp(xy), npts.value, op, C.byref(npts)))
        return xy, offsets

    def solve_packed(self, model, params, keep_dirty=False, xy=None):
        """Solve and return (xy, offsets): all stored polylines densely packed, copied to the host while
        later chunks of the batch are still being solved (dcegm_solve_*_packed)."""
        arr = params if isinstance(params, C.Array) else params_array(params)
        fn = {_abi.MODEL_F: self.lib.dcegm_solve_f_packed, _abi.MODEL_G: self.lib.dcegm_solve_g_packed}[model]
        lay = _abi.dcegm_layout()
        n = len(arr)
        nstate = arr[0].ny if model == _abi.MODEL_G else 1
        nsl = n * arr[0].nT * nstate * 4
        na = arr[0].na
        cap = (na + na // 10 + 2 * max(na // 16, 32) + 2 + 7) // 8 * 8
        if xy is None:
            xy = np.empty((nsl * cap, 2))
        offsets = np.zeros(nsl + 1, np.int64)
        npts = C.c_int64()
        flags = _abi.KEEP_DIRTY if keep_dirty else 0
        self._ck(fn(self.h, arr, n, flags, _dp(xy), xy.shape[0], offsets.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(npts)))
        return xy[: npts.value], offsets

    def fetch_compact(self, first=0, count=None):
        """models [first, first + count) of the resident solution in the compact line format -> CompactLines"""
        from .compact import CompactLines

        lay = self.layout()
        count = lay.nmodels - first if count is None else count
        nsl = count * lay.nT * lay.nstate * 4
        woff = np.zeros(nsl + 1, np.int64)
        lens = np.zeros(nsl, np.int32)
        aux = np.zeros(nsl, np.int32)
        nw = C.c_int64()
        lp, ip = C.POINTER(C.c_int64), C.POINTER(C.c_int32)
        args = (woff.ctypes.data_as(lp), lens.ctypes.data_as(ip), aux.ctypes.data_as(ip), C.byref(nw))
        self._ck(self.lib.dcegm_fetch_compact(self.h, first, count, None, 0, *args))
        words = np.zeros(max(nw.value, 1))
        self._ck(self.lib.dcegm_fetch_compact(self.h, first, count, _dp(words), len(words), *args))
        return CompactLines(lay.nT, lay.nstate, count, words[: nw.value], woff, lens, aux, solver=self, first=first)

    def solve_compact(self, model, params, keep_dirty=False, words=None):
        """Solve and deliver every stored polyline in the compact line format through the chunked pipeline
        (dcegm_solve_*_compact) -> CompactLines"""
        from .compact import CompactLines

        arr = params if isinstance(params, C.Array) else params_array(params)
        fn = {_abi.MODEL_F: self.lib.dcegm_solve_f_compact, _abi.MODEL_G: self.lib.dcegm_solve_g_compact}[model]
        n = len(arr)
        nstate = arr[0].ny if model == _abi.MODEL_G else 1
        nsl = n * arr[0].nT * nstate * 4
        na = arr[0].na
        cap = (na + na // 10 + 2 * max(na // 16, 32) + 2 + 7) // 8 * 8
        if words is None:
            words = np.empty(nsl * (cap + cap // 4) + 1024)  # typical records are ~1.02 n words; E_CAPACITY otherwise
        woff = np.zeros(nsl + 1, np.int64)
        lens = np.zeros(nsl, np.int32)
        aux = np.zeros(nsl, np.int32)
        nw = C.c_int64()
        lp, ip = C.POINTER(C.c_int64), C.POINTER(C.c_int32)
        flags = _abi.KEEP_DIRTY if keep_dirty else 0
        self._ck(fn(self.h, arr, n, flags, _dp(words), len(words), woff.ctypes.data_as(lp), lens.ctypes.data_as(ip),
                    aux.ctypes.data_as(ip), C.byref(nw)))
        return CompactLines(arr[0].nT, nstate, n, words[: nw.value], woff, lens, aux, solver=self, first=0)

    # ---- forward simulation of the resident solution (simulate.jl:61-203) -----------------------
    def simulate(self, panel=False, shocks=None):
        """shocks: None (counter-based draws keyed by Param.rseed) or a dict of the reference's raw rand()
        matrices, see _abi.make_sim_shocks."""
        sp, keep = None, None
        if shocks is not None:
            st, keep = _abi.make_sim_shocks(shocks, self.layout().nT)
            sp = C.byref(st)
        self._ck(self.lib.dcegm_simulate(self.h, _abi.SIM_PANEL if panel else 0, sp))
        del keep

    def sim_moments(self, first=0, count=None):
        lay = self.layout()
        count = lay.nmodels - first if count is None else count
        mom = np.zeros((count, lay.nT, _abi.SIM_NMOM))
        hist = np.zeros((count, lay.nT + 1), np.int32)
        st = np.zeros(count, np.int32)
        ip = C.POINTER(C.c_int32)
        self._ck(self.lib.dcegm_fetch_sim_moments(self.h, first, count, _dp(mom), hist.ctypes.data_as(ip), st.ctypes.data_as(ip)))
        return mom, hist, st

    def sim_moments_to_device(self, ptr, nbytes):
        self._ck(self.lib.dcegm_sim_moments_to_device(self.h, C.c_void_p(ptr), nbytes))

    def sim_panel(self, model, nsims):
        lay = self.layout()
        d = {k: np.zeros((nsims, lay.nT), order="F") for k in ("w0", "w1", "cons", "util", "inc", "prob_work")}
        i = {k: np.zeros((nsims, lay.nT), np.int32, order="F") for k in ("ystate", "worker")}
        ret = np.zeros(nsims, np.int32)
        ip = C.POINTER(C.c_int32)
        self._ck(self.lib.dcegm_fetch_sim_panel(self.h, model, *[_dp(d[k]) for k in ("w0", "w1", "cons", "util", "inc", "prob_work")],
                                                i["ystate"].ctypes.data_as(ip), i["worker"].ctypes.data_as(ip), ret.ctypes.data_as(ip)))
        d.update(i)
        d["ret_age"] = ret
        return d

    def status(self):
        lay = self.layout()
        st = np.zeros(lay.nmodels, np.int32)
        self._ck(self.lib.dcegm_status(self.h, st.ctypes.data_as(C.POINTER(C.c_int32))))
        return st

    def summaries(self):
        lay = self.layout()
        arr = (_abi.dcegm_summary * lay.nmodels)()
        self._ck(self.lib.dcegm_summaries(self.h, arr))
        return np.frombuffer(arr, dtype=SUMMARY_DTYPE).copy()

    def sweep_records_to_device(self, ptr, nbytes, nT_pad):
        """one fixed-size record per model (summary + simulation status + moments padded to nT_pad periods) into
        device memory: the send buffer of the one all-gather of a sweep (dcegm_sweep_records_to_device)"""
        self._ck(self.lib.dcegm_sweep_records_to_device(self.h, C.c_void_p(ptr), nbytes, int(nT_pad)))

    def summaries_to_device(self, ptr, nbytes):
        self._ck(self.lib.dcegm_summaries_to_device(self.h, C.c_void_p(ptr), nbytes))


SUMMARY_DTYPE = np.dtype([
    ("status", np.int32), ("n_env_periods", np.int32), ("tot_segments", np.int32), ("tot_isects", np.int32),
    ("tot_removed", np.int32), ("max_len", np.int32), ("tot_points", np.int64),
    ("v1_probe", np.float64, (_abi.SUMMARY_NPROBE,)), ("c1_probe", np.float64, (_abi.SUMMARY_NPROBE,)),
    ("checksum_c", np.float64), ("checksum_v", np.float64)])
assert SUMMARY_DTYPE.itemsize == C.sizeof(_abi.dcegm_summary)
