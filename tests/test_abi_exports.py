"""The C-ABI library loads on a machine without a GPU and exports every function include/dcegm.h declares;
struct sizes of the ctypes mirror match the header's layout; without a device dcegm_create fails loudly."""
import ctypes as C
import os
import re

import pytest

from dcegm_jl_b200 import _abi, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "dcegm.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dcegm_[a-z_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    lib = _lib.load()
    names = header_functions()
    assert len(names) >= 28
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dcegm.h but not exported"
    assert set(_lib.EXPORTS) <= set(names)
    assert lib.dcegm_abi_version() == _abi.ABI_VERSION


def test_param_defaults_roundtrip():
    lib = _lib.load()
    p = _abi.dcegm_param()
    lib.dcegm_param_defaults(C.byref(p))
    from dcegm_jl_b200 import Param

    q = Param().to_c()
    for n, _ in _abi.dcegm_param._fields_:
        assert getattr(p, n) == getattr(q, n), n


def test_no_device_means_error_not_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.dcegm_create(0, C.byref(h)) == _abi.E_NODEVICE
    assert not h.value
