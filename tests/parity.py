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


def _stored_mask(sol):
    k = np.arange(sol.lay.cap)
    m = k[None, None, None, None, None, :] < sol.lens[..., None]
    return np.broadcast_to(m[..., None], sol.lines.shape)


def assert_bitwise(gpu, ref):
    """GPU vs the oracle in portable-math mode: every stored number identical (NaN == NaN)."""
    assert_same_structure(gpu, ref)
    m = _stored_mask(ref)
    a, b = gpu.lines[m], ref.lines[m]
    same = (a == b) | (np.isnan(a) & np.isnan(b))
    assert same.all(), f"{(~same).sum()} of {same.size} stored ordinates differ; max abs diff {np.nanmax(np.abs(a - b)):.3e}"
    assert np.array_equal(gpu.isects, ref.isects)
    assert np.array_equal(gpu.vbound, ref.vbound)
    if ref.dirty is not None and gpu.dirty is not None:
        assert np.array_equal(gpu.dirty, ref.dirty)


def libm_sensitivity(a, b):
    """Relative deviations between two solutions that differ only in the libm used (1-ulp level
    differences in pow/exp/log).  Returns (share of ordinates within RTOL, max rel. deviation of policy
    kink insertions, max rel. deviation elsewhere).  Requires equal structure."""
    m = _stored_mask(a)
    rel = np.abs(a.lines - b.lines) / (ATOL / RTOL + np.abs(b.lines))
    rel = np.where(m, np.nan_to_num(rel, nan=0.0), 0.0)
    # kink insertions: policy points (kind 0, worker id 0) whose x lies within 1e-9 of an intersection
    kink = np.zeros(a.lines.shape[:-1], bool)
    ni = a.meta[..., 1]
    for (mm, it, iy) in np.argwhere(ni > 0):
        xs = a.isects[mm, it, iy, : ni[mm, it, iy], 0]
        cx = a.lines[mm, it, iy, 0, 0, :, 0]
        near = np.abs(cx[:, None] - xs[None, :]) <= 1e-9 * np.maximum(1.0, np.abs(xs[None, :]))
        kink[mm, it, iy, 0, 0] = near.any(axis=1)
    kink = np.broadcast_to(kink[..., None], rel.shape)
    share = float((rel[m] <= RTOL).mean())
    worst_kink = float(rel[kink & m].max()) if (kink & m).any() else 0.0
    worst_rest = float(rel[~kink & m].max())
    return share, worst_kink, worst_rest
