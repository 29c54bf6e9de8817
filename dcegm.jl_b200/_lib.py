"""Loads csrc/libdcegm_cuda.so and declares the C ABI of include/dcegm.h for ctypes.

There is no CPU fallback: if the library is missing or no CUDA device is present the calls
below raise — nothing in this package computes a solution on the host."""
import ctypes as C
import os

from . import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
# DCEGM_CUDA_LIB selects another build of the same library (kernel A/B experiments, profiling build)
LIB_PATH = os.environ.get("DCEGM_CUDA_LIB") or os.path.join(_HERE, "csrc", "libdcegm_cuda.so")
_lib = None

EXPORTS = [
    "dcegm_abi_version", "dcegm_param_defaults", "dcegm_create", "dcegm_destroy", "dcegm_last_error",
    "dcegm_host_alloc", "dcegm_host_free", "dcegm_solve_f", "dcegm_solve_g", "dcegm_solve_minimal",
    "dcegm_get_layout", "dcegm_fetch_line", "dcegm_fetch_lines", "dcegm_fetch_lens", "dcegm_fetch_vbound",
    "dcegm_fetch_meta", "dcegm_fetch_isects", "dcegm_fetch_removed", "dcegm_fetch_dirty", "dcegm_fetch_packed",
    "dcegm_status", "dcegm_summaries", "dcegm_summaries_to_device", "dcegm_last_timing", "dcegm_upload_f",
    "dcegm_upload_g", "dcegm_upload_minimal", "dcegm_run", "dcegm_debug_math_check", "dcegm_debug_dfma_rate", "dcegm_solve_f_packed", "dcegm_solve_g_packed",
    "dcegm_simulate", "dcegm_fetch_sim_moments", "dcegm_fetch_sim_panel", "dcegm_sim_moments_to_device", "dcegm_solve_minimal_borrow", "dcegm_solve_minimal_bequest",
    "dcegm_sweep_record_bytes", "dcegm_sweep_records_to_device",
    "dcegm_solve_f_compact", "dcegm_solve_g_compact", "dcegm_fetch_compact", "dcegm_expand_compact",
    "dcegm_decode_compact_pair", "dcegm_get_grid", "dcegm_debug_host_table", "dcegm_debug_check_guards",
    "dcegm_pack_lines_noalloc",
]


class DcegmError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"libdcegm_cuda error {code}: {text}")
        self.code = code


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `make -C dcegm.jl_b200/csrc` (or __graft_entry__.build()). "
            "This package has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    H = C.c_void_p
    PP = C.POINTER(_abi.dcegm_param)
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int32)
    lib.dcegm_abi_version.restype = C.c_int
    lib.dcegm_param_defaults.argtypes = [PP]
    lib.dcegm_param_defaults.restype = None
    lib.dcegm_create.argtypes = [C.c_int, C.POINTER(H)]
    lib.dcegm_destroy.argtypes = [H]
    lib.dcegm_last_error.argtypes = [H]
    lib.dcegm_last_error.restype = C.c_char_p
    lib.dcegm_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
    lib.dcegm_host_free.argtypes = [C.c_void_p]
    for f in ("dcegm_solve_f", "dcegm_solve_g", "dcegm_solve_minimal", "dcegm_solve_minimal_borrow", "dcegm_solve_minimal_bequest", "dcegm_upload_f", "dcegm_upload_g",
              "dcegm_upload_minimal"):
        getattr(lib, f).argtypes = [H, PP, C.c_int, C.c_uint]
    for f in ("dcegm_solve_f_packed", "dcegm_solve_g_packed"):
        getattr(lib, f).argtypes = [H, PP, C.c_int, C.c_uint, dp, C.c_size_t, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lp = C.POINTER(C.c_int64)
    for f in ("dcegm_solve_f_compact", "dcegm_solve_g_compact"):
        getattr(lib, f).argtypes = [H, PP, C.c_int, C.c_uint, dp, C.c_size_t, lp, ip, ip, lp]
    lib.dcegm_fetch_compact.argtypes = [H, C.c_int, C.c_int, dp, C.c_size_t, lp, ip, ip, lp]
    lib.dcegm_expand_compact.argtypes = [H, C.c_int, C.c_int, dp, lp, ip, ip, dp, C.c_size_t, lp, lp, C.c_int]
    lib.dcegm_decode_compact_pair.argtypes = [dp, C.c_int, dp, C.c_int, C.c_int, dp, C.c_int, C.c_int, dp, dp]
    lib.dcegm_get_grid.argtypes = [H, C.c_int, dp, C.c_int, C.POINTER(C.c_int)]
    lib.dcegm_debug_check_guards.argtypes = [H, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    # device pointers travel as plain integers (torch data_ptr())
    lib.dcegm_pack_lines_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_size_t,
                                             C.c_void_p, C.POINTER(C.c_int64), C.c_void_p, C.c_char_p, C.c_size_t]
    lib.dcegm_debug_host_table.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, dp, dp]
    lib.dcegm_simulate.argtypes = [H, C.c_uint, C.POINTER(_abi.dcegm_sim_shocks)]
    lib.dcegm_fetch_sim_moments.argtypes = [H, C.c_int, C.c_int, dp, ip, ip]
    lib.dcegm_fetch_sim_panel.argtypes = [H, C.c_int, dp, dp, dp, dp, dp, dp, ip, ip, ip]
    lib.dcegm_sim_moments_to_device.argtypes = [H, C.c_void_p, C.c_size_t]
    lib.dcegm_sweep_record_bytes.argtypes = [C.c_int]
    lib.dcegm_sweep_records_to_device.argtypes = [H, C.c_void_p, C.c_size_t, C.c_int]
    lib.dcegm_run.argtypes = [H]
    lib.dcegm_get_layout.argtypes = [H, C.POINTER(_abi.dcegm_layout)]
    lib.dcegm_fetch_line.argtypes = [H, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, dp, C.c_int, C.POINTER(C.c_int)]
    lib.dcegm_fetch_lines.argtypes = [H, C.c_int, C.c_int, dp]
    lib.dcegm_fetch_lens.argtypes = [H, C.c_int, C.c_int, ip]
    lib.dcegm_fetch_vbound.argtypes = [H, C.c_int, C.c_int, dp]
    lib.dcegm_fetch_meta.argtypes = [H, C.c_int, C.c_int, C.c_void_p]
    lib.dcegm_fetch_isects.argtypes = [H, C.c_int, C.c_int, dp]
    lib.dcegm_fetch_removed.argtypes = [H, C.c_int, C.c_int, C.POINTER(C.c_uint32)]
    lib.dcegm_fetch_dirty.argtypes = [H, C.c_int, C.c_int, dp]
    lib.dcegm_fetch_packed.argtypes = [H, C.c_int, C.c_int, dp, C.c_size_t, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.dcegm_status.argtypes = [H, ip]
    lib.dcegm_summaries.argtypes = [H, C.POINTER(_abi.dcegm_summary)]
    lib.dcegm_summaries_to_device.argtypes = [H, C.c_void_p, C.c_size_t]
    lib.dcegm_last_timing.argtypes = [H, C.POINTER(_abi.dcegm_timing)]
    lib.dcegm_debug_dfma_rate.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.dcegm_debug_math_check.argtypes = [C.c_ulonglong, C.c_ulonglong, C.c_int, C.c_int, C.POINTER(C.c_ulonglong)]
    for name in EXPORTS:
        fn = getattr(lib, name)  # AttributeError here means the library does not match include/dcegm.h
        if name == "dcegm_sweep_record_bytes":
            fn.restype = C.c_size_t
        elif name not in ("dcegm_last_error", "dcegm_param_defaults"):
            fn.restype = C.c_int
    if lib.dcegm_abi_version() != _abi.ABI_VERSION:
        raise RuntimeError("libdcegm_cuda.so ABI version mismatch")
    _lib = lib
    return lib
