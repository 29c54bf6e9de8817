"""ctypes mirror of include/dcegm.h (structures and constants only; no library is loaded here)."""
import ctypes as C

ABI_VERSION = 1

# error codes
OK, E_INVALID, E_CUDA, E_NOSOLUTION, E_NODEVICE, E_CAPACITY = 0, -1, -2, -3, -4, -5

# per-model status bits
ST_ISECT_OVERFLOW = 1 << 0
ST_LINE_OVERFLOW = 1 << 1
ST_NONMONOTONE = 1 << 2
ST_NO_BRACKET = 1 << 3
ST_MISALIGNED = 1 << 4
ST_ALL_FALLING = 1 << 5
ST_NONFINITE = 1 << 6
ST_SEG_OVERFLOW = 1 << 7
ST_INSERT_RANGE = 1 << 8
ST_LOOP_LIMIT = 1 << 9

STATUS_TEXT = {
    ST_ISECT_OVERFLOW: "more intersections than isec_cap",
    ST_LINE_OVERFLOW: "a line outgrew its slot",
    ST_NONMONOTONE: "knot vector not strictly increasing (Interpolations.jl would throw, line.jl:236)",
    ST_NO_BRACKET: "fzero without a sign change (envelope.jl:199)",
    ST_MISALIGNED: "insert_left/insert_right lengths differ (dc_algo.jl:486-491)",
    ST_ALL_FALLING: "splitLine produced no section (envelope.jl:97-104)",
    ST_NONFINITE: "NaN/Inf in a computed line",
    ST_SEG_OVERFLOW: "more sections than seg_cap",
    ST_INSERT_RANGE: "kink insertion index out of range (dc_algo.jl:472,489)",
    ST_LOOP_LIMIT: "third-line loop did not terminate (envelope.jl:197-240)",
}

MODEL_F, MODEL_G, MODEL_MINIMAL = 1, 2, 3
KIND_C, KIND_V, KIND_CDIRTY, KIND_VDIRTY = 0, 1, 2, 3
KEEP_DIRTY = 1 << 0
SUMMARY_ONLY = 1 << 1
SUMMARY_NPROBE = 8

_DOUBLES = [
    "gamma", "beta", "sigma", "lambda_", "R", "a_high", "a_low", "a_lowT", "cfloor", "cfloor_plot",
    "alpha", "alphaT", "inc0", "inc1", "inc2", "rho", "pension", "delta", "maxinterest", "bbar", "nu",
    "initw0", "initw1",
]
_INTS = ["na", "ny", "nT", "nD", "k", "retage", "nsims", "quad_group"]


class dcegm_param(C.Structure):
    _fields_ = [(n, C.c_double) for n in _DOUBLES] + [(n, C.c_int32) for n in _INTS] + [("rseed", C.c_int64)]


class dcegm_layout(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "model", "nmodels", "na", "ny", "nT", "nstate", "cap", "isec_cap", "rem_words", "seg_cap",
        "has_dirty", "has_lines")]


class dcegm_env_meta(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_segments", "n_isects", "n_removed", "n_prepended")]


class dcegm_summary(C.Structure):
    _fields_ = [
        ("status", C.c_int32), ("n_env_periods", C.c_int32), ("tot_segments", C.c_int32),
        ("tot_isects", C.c_int32), ("tot_removed", C.c_int32), ("max_len", C.c_int32),
        ("tot_points", C.c_int64),
        ("v1_probe", C.c_double * SUMMARY_NPROBE), ("c1_probe", C.c_double * SUMMARY_NPROBE),
        ("checksum_c", C.c_double), ("checksum_v", C.c_double),
    ]


class dcegm_timing(C.Structure):
    _fields_ = [("setup_host_ms", C.c_float), ("h2d_ms", C.c_float), ("kernel_ms", C.c_float),
                ("n_launches", C.c_int32), ("_pad", C.c_int32)]


assert C.sizeof(dcegm_param) == 23 * 8 + 8 * 4 + 8
assert C.sizeof(dcegm_summary) == 6 * 4 + 8 + 2 * 8 * SUMMARY_NPROBE + 16
