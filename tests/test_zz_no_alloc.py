"""dcegm_pack_lines_noalloc (include/dcegm.h, csrc/no_alloc.cu): the packing of dcegm_fetch_packed on caller-provided
device memory, with its offsets scratch in a static device array — no allocation inside the call.  Checked against
dcegm_fetch_packed BIT FOR BIT on two calls in a row (the second on another solution: the shared array keeps the
first call's contents), plus the size limit, the capacity error and the absence of allocations.

Collected last on purpose: a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_SLOTS = 1 << 20  # DCEGM_NOALLOC_MAX_SLOTS


def retirement_params(n, na):
    from dcegm_jl_b200 import Param

    rng = np.random.default_rng(25032020)
    lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), size=64)
    sig = rng.uniform(0, 0.6, size=64)
    ps = []
    for i in range(n):
        d = dict(na=na, beta=0.95, R=1.05, sigma=float(sig[i]))
        d["lambda"] = float(lam[i])
        ps.append(Param(par=d))
    return ps


def pack_noalloc(lib, torch, lines, lens, cap_points=None, want_total=True, stream=None):
    """lines/lens: host arrays of a Solution -> (rc, message, xy, offsets, npoints) through the new entry point"""
    dev = torch.device("cuda", 0)
    cap = lines.shape[-2]
    nsl = lens.size
    d_lines = torch.from_numpy(np.ascontiguousarray(lines)).to(dev)
    d_lens = torch.from_numpy(np.ascontiguousarray(lens)).to(dev)
    room = int(np.minimum(lens, cap).sum()) if cap_points is None else cap_points
    d_xy = torch.full((max(room, 1) + 8, 2), -7.0, dtype=torch.float64, device=dev)  # 8 points of canary behind the room
    d_off = torch.full((nsl + 1,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    npts = C.c_int64(-1)
    err = C.create_string_buffer(256)
    rc = lib.dcegm_pack_lines_noalloc(0, d_lines.data_ptr(), d_lens.data_ptr(), nsl, cap, d_xy.data_ptr(), room,
                                      d_off.data_ptr(), C.byref(npts) if want_total else None,
                                      stream.cuda_stream if stream is not None else None, err, len(err))
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info(0)[0]
    return rc, err.value.decode(), d_xy.cpu().numpy(), d_off.cpu().numpy(), npts.value, room, free0 - free1


def test_noalloc_packing_equals_fetch_packed_twice():
    import torch

    from dcegm_jl_b200 import Param, _abi
    from dcegm_jl_b200.solver import Solver

    with Solver(0) as s:
        lib = s.lib
        # call 1: retirement models (envelope-heavy draws: lines of very different lengths, empty slots)
        sol = s.solve(_abi.MODEL_F, retirement_params(24, 300))
        xy_ref, off_ref = s.fetch_packed()
        rc, msg, xy, off, npts, room, _ = pack_noalloc(lib, torch, sol.lines, sol.lens)
        assert rc == 0, msg
        assert npts == off_ref[-1] == room
        assert np.array_equal(off, off_ref)
        assert xy[:room].tobytes() == xy_ref.tobytes()
        assert (xy[room:] == -7.0).all()
        # call 2, right behind it: general models, other geometry — nothing of call 1 may show through
        pg = [Param(par=dict(na=100, rho=r, sigma=sg)) for r, sg in ((0.9, 0.1), (0.3, 0.2), (0.0, 0.3))]
        sol_g = s.solve(_abi.MODEL_G, pg)
        xy_ref_g, off_ref_g = s.fetch_packed()
        rc, msg, xy_g, off_g, npts_g, room_g, used = pack_noalloc(lib, torch, sol_g.lines, sol_g.lens)
        assert rc == 0, msg
        assert used == 0, f"device memory in use grew by {used} bytes across the call"
        assert npts_g == off_ref_g[-1] == room_g
        assert np.array_equal(off_g, off_ref_g)
        assert xy_g[:room_g].tobytes() == xy_ref_g.tobytes()
        # call 3: the first input again, asynchronous form (no host total) on a side stream
        st = torch.cuda.Stream(device=0)
        rc, msg, xy3, off3, npts3, _, used = pack_noalloc(lib, torch, sol.lines, sol.lens, want_total=False, stream=st)
        assert rc == 0 and npts3 == -1 and used == 0
        assert np.array_equal(off3, off_ref) and xy3[:room].tobytes() == xy_ref.tobytes()


def test_noalloc_limits_and_capacity():
    import torch

    from dcegm_jl_b200 import _abi
    from dcegm_jl_b200.solver import Solver

    with Solver(0) as s:
        lib = s.lib
        sol = s.solve(_abi.MODEL_F, retirement_params(4, 200))
        xy_ref, off_ref = s.fetch_packed()
        total = int(off_ref[-1])
        # destination too small: error code and message, total still reported, nothing written behind the room,
        # every slot that fits is in place
        small = total - 5
        rc, msg, xy, off, npts, room, _ = pack_noalloc(lib, torch, sol.lines, sol.lens, cap_points=small)
        assert rc == _abi.E_CAPACITY and "too small" in msg
        assert npts == total and np.array_equal(off, off_ref)
        assert (xy[small:] == -7.0).all()
        fits = int(np.searchsorted(off_ref, small, side="right")) - 1  # slots [0, fits) end within the room
        assert xy[: off_ref[fits]].tobytes() == xy_ref[: off_ref[fits]].tobytes()
        # ... and the next call is clean again (the overflow count was cleared)
        rc, msg, xy, off, npts, room, _ = pack_noalloc(lib, torch, sol.lines, sol.lens)
        assert rc == 0 and xy[:room].tobytes() == xy_ref.tobytes()
        # beyond the static scratch: rejected before any device work, with the limit in the message
        err = C.create_string_buffer(256)
        d = torch.zeros(16, dtype=torch.float64, device="cuda:0")
        rc = lib.dcegm_pack_lines_noalloc(0, d.data_ptr(), d.data_ptr(), MAX_SLOTS + 1, 4, d.data_ptr(), 8, None, None, None,
                                          err, len(err))
        assert rc == _abi.E_INVALID
        assert "DCEGM_NOALLOC_MAX_SLOTS" in err.value.decode() and str(MAX_SLOTS) in err.value.decode()
        # null pointers and an empty input
        assert lib.dcegm_pack_lines_noalloc(0, None, d.data_ptr(), 1, 4, d.data_ptr(), 8, None, None, None, err, len(err)) == _abi.E_INVALID
        n0 = C.c_int64(-1)
        assert lib.dcegm_pack_lines_noalloc(0, d.data_ptr(), d.data_ptr(), 0, 4, d.data_ptr(), 8, None, C.byref(n0), None, err, len(err)) == 0
        assert n0.value == 0
