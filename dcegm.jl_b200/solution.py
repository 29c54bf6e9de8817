"""Host-side view of a batch solution in the flat layout documented in include/dcegm.h."""
import numpy as np

from . import _abi


class Solution:
    """lines  [nmodels, nT, nstate, 2, 2, cap, 2]   lens [nmodels, nT, nstate, 2, 2]
    vbound [nmodels, nT, nstate, 2]   meta [nmodels, nT, nstate, 4]
    isects [nmodels, nT, nstate, isec_cap, 2]   removed [nmodels, nT, nstate, rem_words] (uint32 bitmask)
    dirty  [nmodels, nT, nstate, 2, 2, na, 2] or None   status [nmodels]"""

    def __init__(self, lay, lines, lens, vbound, meta, isects, removed, dirty, status):
        self.lay = lay
        self.lines, self.lens, self.vbound, self.meta = lines, lens, vbound, meta
        self.isects, self.removed, self.dirty, self.status = isects, removed, dirty, status

    @staticmethod
    def shapes(lay):
        nm, nT, ns = lay.nmodels, lay.nT, lay.nstate
        return dict(
            lines=(nm, nT, ns, 2, 2, lay.cap, 2), lens=(nm, nT, ns, 2, 2), vbound=(nm, nT, ns, 2),
            meta=(nm, nT, ns, 4), isects=(nm, nT, ns, lay.isec_cap, 2), removed=(nm, nT, ns, lay.rem_words),
            dirty=(nm, nT, ns, 2, 2, lay.na, 2), status=(nm,),
        )

    @classmethod
    def allocate(cls, lay, keep_dirty=False):
        s = cls.shapes(lay)
        return cls(
            lay, np.zeros(s["lines"]), np.zeros(s["lens"], np.int32), np.zeros(s["vbound"]),
            np.zeros(s["meta"], np.int32), np.zeros(s["isects"]), np.zeros(s["removed"], np.uint32),
            np.zeros(s["dirty"]) if keep_dirty else None, np.zeros(s["status"], np.int32),
        )

    # accessors (0-based it / id / iy) -------------------------------------------------
    def line(self, model, kind, id, it, iy=0):
        if kind in (_abi.KIND_CDIRTY, _abi.KIND_VDIRTY):
            a = self.dirty[model, it, iy, id, kind - 2]
            return a[:, 0].copy(), a[:, 1].copy()
        n = int(self.lens[model, it, iy, id, kind])
        a = self.lines[model, it, iy, id, kind, :n]
        return a[:, 0].copy(), a[:, 1].copy()

    def length(self, model, kind, id, it, iy=0):
        return int(self.lens[model, it, iy, id, kind])

    def isect_points(self, model, it, iy=0):
        n = int(self.meta[model, it, iy, 1])
        return self.isects[model, it, iy, :n].copy()

    def removed_idx(self, model, it, iy=0):
        """0-based indices into the line handed to secondary_envelope (incl. bend-back points)"""
        bits = np.unpackbits(self.removed[model, it, iy].view(np.uint8), bitorder="little")
        return np.nonzero(bits)[0]

    def status_text(self, model):
        st = int(self.status[model])
        return [t for b, t in _abi.STATUS_TEXT.items() if st & b]
