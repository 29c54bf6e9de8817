"""ctypes mirror of include/dcegm.h (structures and constants only; no library is loaded here)."""
import ctypes as C

ABI_VERSION = 2

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
ST_NEG_CONS = 1 << 10
ST_YSTATE_RANGE = 1 << 11
# informational: the reference returns normally on these (include/dcegm.h)
ST_EXTRA_RIGHT = 1 << 12
ST_NONMONO_FIRST = 1 << 13
ST_DUP_KNOT = 1 << 14
ST_INFO_MASK = ST_EXTRA_RIGHT | ST_NONMONO_FIRST
ST_FATAL_MASK = ~ST_INFO_MASK & 0xFFFFFFFF

STATUS_TEXT = {
    ST_ISECT_OVERFLOW: "more intersections than isec_cap",
    ST_LINE_OVERFLOW: "a line outgrew its slot",
    ST_NONMONOTONE: "knot vector not strictly increasing (Interpolations.jl would throw, line.jl:236)",
    ST_NO_BRACKET: "fzero without a sign change (envelope.jl:199)",
    ST_MISALIGNED: "insert_right shorter than insert_left (dc_algo.jl:486-491)",
    ST_ALL_FALLING: "splitLine produced no section (envelope.jl:97-104)",
    ST_NONFINITE: "NaN/Inf in a computed line",
    ST_SEG_OVERFLOW: "more sections than seg_cap",
    ST_INSERT_RANGE: "kink insertion index out of range (dc_algo.jl:472,489)",
    ST_LOOP_LIMIT: "third-line loop did not terminate (envelope.jl:197-240)",
    ST_NEG_CONS: "simulator: negative consumption (simulate.jl:111-116 calls error())",
    ST_YSTATE_RANGE: "simulator: Markov draw beyond the last cumulative weight (simulate.jl:148)",
    ST_DUP_KNOT: "a stored line repeats an abscissa (intersection at a shared turning point, envelope.jl:199,212,248-252)",
    ST_EXTRA_RIGHT: "info: insert_right longer than insert_left, surplus dropped (dc_algo.jl:486-492)",
    ST_NONMONO_FIRST: "info: a period-1 line is not strictly increasing (never interpolated by the solve)",
}

MODEL_F, MODEL_G, MODEL_MINIMAL = 1, 2, 3
MODEL_MINIMAL_BORROW, MODEL_MINIMAL_BEQUEST = 4, 5
KIND_C, KIND_V, KIND_CDIRTY, KIND_VDIRTY = 0, 1, 2, 3
KEEP_DIRTY = 1 << 0
SUMMARY_ONLY = 1 << 1
FORCE_FINE_GRID = 1 << 2
SUMMARY_NPROBE = 8
SIM_PANEL = 1 << 0
SIM_NMOM = 8
SIM_MOMENTS = ("w0", "w1", "cons", "util", "inc", "prob_work", "share_worker", "share_neg_cons")

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


class dcegm_sim_shocks(C.Structure):
    _fields_ = [("wshocks", C.POINTER(C.c_double)), ("dshocks", C.POINTER(C.c_double)), ("rshocks", C.POINTER(C.c_double)),
                ("w0shocks", C.POINTER(C.c_double)), ("ystate0", C.POINTER(C.c_int32)), ("nsims", C.c_int32), ("_pad", C.c_int32)]


def make_sim_shocks(shocks, nT):
    """dict(wshocks, dshocks, w0shocks[, rshocks][, ystate0]) of the reference's raw rand() draws, matrices
    (nsims, nT) -> (struct, keep-alive list).  Matrices are laid out column-major like Julia's."""
    import numpy as np

    keep = []

    def mat(name):
        a = shocks.get(name)
        if a is None:
            return None
        a = np.asfortranarray(np.asarray(a, float))
        keep.append(a)
        return a.ctypes.data_as(C.POINTER(C.c_double))

    s = dcegm_sim_shocks()
    w0 = np.ascontiguousarray(shocks["w0shocks"], float)
    keep.append(w0)
    s.nsims = len(w0)
    assert np.asarray(shocks["wshocks"]).shape == (s.nsims, nT)
    s.wshocks, s.dshocks, s.rshocks = mat("wshocks"), mat("dshocks"), mat("rshocks")
    s.w0shocks = w0.ctypes.data_as(C.POINTER(C.c_double))
    if shocks.get("ystate0") is not None:
        y0 = np.ascontiguousarray(shocks["ystate0"], np.int32)
        keep.append(y0)
        s.ystate0 = y0.ctypes.data_as(C.POINTER(C.c_int32))
    return s, keep
