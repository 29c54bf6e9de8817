"""dcegm.jl_b200 — B200-native DC-EGM solver behind DCEGM.jl's entry points.

Host-side mirror (Python, because no Julia toolchain exists in this image) of the reference's
public interface for the hot path: ``Param``, ``runf``/``runfdemo``/``rung``/``bm``/``minimal_EGM``
and the ``MLine``/``Envelope`` result containers (api.py).  All numerical work happens in
``csrc/libdcegm_cuda.so`` (hand-written sm_100a CUDA behind the C ABI of include/dcegm.h);
there is no CPU fallback — importing works anywhere, solving requires the library and a GPU.
"""
from . import _abi
from .param import Param, DEFAULTS
from .solution import Solution
from .api import (Point, MLine, Envelope, FModel, GModel, runf, runfdemo, rung, minimal_EGM, bm, getx, gety, gets,
                  getr, bound, DCEGMError, Simulation, sim, rs, rsg, minimal_EGM_borrow, minimal_EGM_bequest, v_analytic)
from . import io

__all__ = ["Param", "DEFAULTS", "Solution", "_abi", "Point", "MLine", "Envelope", "FModel", "GModel", "runf",
           "runfdemo", "rung", "minimal_EGM", "bm", "getx", "gety", "gets", "getr", "bound", "DCEGMError",
           "Simulation", "sim", "rs", "rsg", "minimal_EGM_borrow", "minimal_EGM_bequest", "v_analytic", "io"]
