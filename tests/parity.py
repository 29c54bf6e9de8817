"""Helpers shared by the GPU parity tests: compare a CUDA Solution with the oracle's."""
import numpy as np

RTOL = 1e-10   # north_star: values/policies within 1e-10 relative (FP64)
ATOL = 1e-12


def line_errors(a, b):
    """max relative error over all stored points of two Solutions with equal lens"""
    lay = a.lay
    k = np.arange(lay.cap)
    mask = k[None, None, None, None, None, :] < a.lens[..., None]
    mask = np.broadcast_to(mask[..., None], a.lines.shape)
    x, y = a.lines[mask], b.lines[mask]
    both_nan = np.isnan(x) & np.isnan(y)
    x, y = x[~both_nan], y[~both_nan]
    if np.isnan(x).any() or np.isnan(y).any():
        return np.inf
    return float(np.max(np.abs(x - y) / (ATOL / RTOL + np.abs(y)))) if x.size else 0.0


def assert_same_structure(gpu, ref):
    assert np.array_equal(gpu.status, ref.status), (gpu.status, ref.status)
    assert np.array_equal(gpu.lens, ref.lens), np.argwhere(gpu.lens != ref.lens)[:10]
    assert np.array_equal(gpu.meta, ref.meta), np.argwhere(gpu.meta != ref.meta)[:10]
    assert np.array_equal(gpu.removed, ref.removed), np.argwhere(gpu.removed != ref.removed)[:10]


def assert_parity(gpu, ref, rtol=RTOL):
    """discrete structure exact; all ordinates within rtol"""
    assert_same_structure(gpu, ref)
    err = line_errors(gpu, ref)
    assert err <= rtol, f"max relative line error {err:.3e} > {rtol:.1e}"
    ni = ref.meta[..., 1]
    k = np.arange(ref.lay.isec_cap)
    m = np.broadcast_to((k[None, None, None, :] < ni[..., None])[..., None], ref.isects.shape)
    if m.any():
        e = np.max(np.abs(gpu.isects[m] - ref.isects[m]) / (ATOL / RTOL + np.abs(ref.isects[m])))
        assert e <= rtol, f"intersection error {e:.3e}"
    vb = np.abs(gpu.vbound - ref.vbound) / (ATOL / RTOL + np.abs(ref.vbound))
    assert vb.max() <= rtol
    return err
