"""CompactLines — host-side view of a batch solution delivered in the compact line format (include/dcegm.h,
"compact lines"): ordinates plus abscissa runs, about half the bytes of Vector{Point} storage.  Lines are rebuilt
lazily, one (model, it, iy, id) pair at a time (`line`), or all at once on host threads (`expand`); both give the
stored points bit for bit.  A Julia mirror would do the same behind `m.c[id,it].env` / `m.v[id,it].env`."""
import ctypes as C
import os

import numpy as np

from . import _abi, _lib

KIND_GRID, KIND_REF, KIND_EXPLICIT = 0, 1, 2
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lp = C.POINTER(C.c_int64)


class CompactLines:
    """words [nwords] float64 (8-byte words), woffsets [nslots + 1], lens [nslots], aux [nslots]; slots in layout
    order [model][it][iy][id][kind].  woffsets[slot] is where the slot's record starts; a slot whose line is bit-identical
    to the same slot of its delivery chunk's first model has that model's offset (no record of its own), so sizes are
    never differences of offsets.  `solver` (optional) lets `expand` use the library's threaded decoder and
    `grid` fetch the savings grids; without it pass `grids` = {model: avec}."""

    def __init__(self, lay_nT, lay_nstate, nmodels, words, woffsets, lens, aux, solver=None, first=0, grids=None):
        self.nT, self.nstate, self.nmodels = lay_nT, lay_nstate, nmodels
        self.words, self.woffsets, self.lens, self.aux = words, woffsets, lens, aux
        self.solver, self.first = solver, first
        self._grids = dict(grids or {})

    @property
    def nbytes(self):
        """bytes delivered to the host: records + per-slot offsets, lengths and run counts"""
        return int(self.woffsets[-1]) * 8 + self.woffsets.nbytes + self.lens.nbytes + self.aux.nbytes

    def slot(self, model, kind, id, it, iy=0):
        return (((model * self.nT + it) * self.nstate + iy) * 2 + id) * 2 + kind

    def grid(self, model):
        g = self._grids.get(model)
        if g is None:
            s = self.solver
            na = C.c_int()
            s._ck(s.lib.dcegm_get_grid(s.h, self.first + model, None, 0, C.byref(na)))
            g = np.zeros(na.value)
            s._ck(s.lib.dcegm_get_grid(s.h, self.first + model, g.ctypes.data_as(_dp), na.value, C.byref(na)))
            self._grids[model] = g
        return g

    def line(self, model, kind, id, it, iy=0):
        """(x, y) of one stored line, decoded by the handle-free primitive dcegm_decode_compact_pair"""
        lib = _lib.load()
        sc = self.slot(model, _abi.KIND_C, id, it, iy)
        sv = sc + 1
        nc, nv = int(self.lens[sc]), int(self.lens[sv])
        a = self.grid(model)
        xy_c = np.zeros((max(nc, 1), 2))
        xy_v = np.zeros((max(nv, 1), 2))
        rc_ = self.words[int(self.woffsets[sc]):]
        rv_ = self.words[int(self.woffsets[sv]):]
        want_v = kind == _abi.KIND_V
        rc = lib.dcegm_decode_compact_pair(a.ctypes.data_as(_dp), len(a), rc_.ctypes.data_as(_dp), nc, int(self.aux[sc]),
                                           rv_.ctypes.data_as(_dp) if want_v else None, nv if want_v else 0,
                                           int(self.aux[sv]) if want_v else 0, xy_c.ctypes.data_as(_dp), xy_v.ctypes.data_as(_dp))
        if rc != 0:
            raise _lib.DcegmError(rc, "inconsistent compact record")
        out, n = (xy_v, nv) if want_v else (xy_c, nc)
        return out[:n, 0].copy(), out[:n, 1].copy()

    def expand(self, nthreads=None):
        """(xy [npoints, 2], offsets [nslots + 1]): the densely packed points dcegm_fetch_packed delivers"""
        s = self.solver
        nsl = len(self.lens)
        offsets = np.zeros(nsl + 1, np.int64)
        npts = C.c_int64()
        total = int(self.lens.sum())
        xy = np.empty((total, 2))
        nthreads = nthreads or (os.cpu_count() or 1)
        s._ck(s.lib.dcegm_expand_compact(s.h, self.first, self.nmodels, self.words.ctypes.data_as(_dp),
                                         self.woffsets.ctypes.data_as(_lp), self.lens.ctypes.data_as(_ip),
                                         self.aux.ctypes.data_as(_ip), xy.ctypes.data_as(_dp), total,
                                         offsets.ctypes.data_as(_lp), C.byref(npts), int(nthreads)))
        return xy, offsets

    def runs(self, slot):
        """[(pos, src, kind)] of one slot (diagnostics, tests)"""
        nrun = int(self.aux[slot]) & 0xFFFF
        w0 = int(self.woffsets[slot])
        if int(self.lens[slot]) == 0:
            return []
        yw = int(self.words[w0: w0 + 1].view(np.uint64)[0])
        e = self.words[w0 + 1 + yw: w0 + 1 + yw + nrun].view(np.uint64)
        return [(int(v & 0xFFFF), int((v >> 16) & 0xFFFF), int((v >> 32) & 0xFFFF)) for v in e]


def encode_reference(lines_c, lines_v, grid):
    """Pure-numpy encoder of ONE (policy, value) pair following the format's rules (tests of the host decoder on a
    machine without a GPU; the product encoder is csrc/compact_lines.cu).  lines_*: (n, 2) arrays.  Returns
    (rec_c, aux_c, rec_v, aux_v)."""

    def enc_y(y):
        """ordinate section: I_0 | I_1 - I_0 | control bytes | payload; blocks of 8 points hold the second or the third
        differences of the bit patterns (whichever is narrower), w four-bit units per residual"""
        n = len(y)
        if n == 0:
            return np.zeros(0)
        M = (1 << 64) - 1
        I = [int(v) for v in np.ascontiguousarray(y, float).view(np.uint64)]
        head = [I[0]] + ([(I[1] - I[0]) & M] if n > 1 else [])
        R2 = [0] * n  # R2_0 = R2_1 = 0 (raw words in front)
        for k in range(2, n):
            R2[k] = (I[k] - 2 * I[k - 1] + I[k - 2]) & M
        R3 = [0] * n
        for k in range(2, n):
            R3[k] = (R2[k] - R2[k - 1]) & M

        def signed(u):
            return u - (1 << 64) if u >> 63 else u

        def nibbles(u):
            r = signed(u)
            return 0 if r == 0 else ((r if r >= 0 else ~r).bit_length() + 1 + 3) // 4

        ctl, payload = [], bytearray()
        for b0 in range(0, n, 8):
            cnt = min(8, n - b0)
            w2 = max(nibbles(u) for u in R2[b0:b0 + cnt])
            w3 = max(nibbles(u) for u in R3[b0:b0 + cnt])
            third = w3 < w2
            w, R = (w3, R3) if third else (w2, R2)
            ctl.append(w | (32 if third else 0))
            hb = w >> 1
            for u in R[b0:b0 + cnt]:
                payload += u.to_bytes(8, "little")[:hb]
            if w & 1:
                tops = [(u >> (8 * hb)) & 15 for u in R[b0:b0 + cnt]] + [0]
                payload += bytes(tops[2 * i] | (tops[2 * i + 1] << 4) for i in range((cnt + 1) // 2))
        cb = bytes(ctl) + bytes(-len(ctl) % 8)
        payload += bytes(-len(payload) % 8)
        return np.concatenate([np.array(head, np.uint64), np.frombuffer(cb + bytes(payload), np.uint64)]).view(np.float64)

    def enc(L, codes):
        n = len(L)
        runs, xs = [], []
        prev = None
        for k, (kind, src) in enumerate(codes):
            new = prev is None or prev[0] != kind or (kind != KIND_EXPLICIT and src != prev[1] + 1)
            if new:
                runs.append(k | (src << 16) | (kind << 32))
            if kind == KIND_EXPLICIT:
                xs.append(L[k, 0])
            prev = (kind, src)
        if len(runs) + len(xs) > n + 1:
            runs, xs = [KIND_EXPLICIT << 32], list(L[:, 0])
        if n == 0:
            return np.zeros(0), 0
        ysec = enc_y(L[:, 1])
        rec = np.concatenate([np.array([len(ysec)], np.uint64).view(np.float64), ysec,
                              np.array(runs, np.uint64).view(np.float64), np.array(xs, float)])
        return rec, len(runs) | (len(xs) << 16)

    cc = []
    for x, y in lines_c:
        hit = [i for i in range(len(grid)) if (np.float64(grid[i]) + np.float64(y)).tobytes() == np.float64(x).tobytes()]
        cc.append((KIND_GRID, hit[0]) if hit else (KIND_EXPLICIT, 0))
    cv = []
    xc = lines_c[:, 0] if len(lines_c) else np.zeros(0)
    for x, _ in lines_v:
        hit = [j for j in range(len(xc)) if np.float64(xc[j]).tobytes() == np.float64(x).tobytes()]
        cv.append((KIND_REF, hit[-1]) if hit else (KIND_EXPLICIT, 0))
    rc, ac = enc(np.asarray(lines_c, float).reshape(-1, 2), cc)
    rv, av = enc(np.asarray(lines_v, float).reshape(-1, 2), cv)
    return rc, ac, rv, av
