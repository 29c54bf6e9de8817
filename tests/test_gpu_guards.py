"""Stand-ins for compute-sanitizer, which is closed on this GPU pool (profiles/r2_experiments.md):
  * memcheck: every device buffer of a handle ends in a guard band written at allocation; after every solver, the
    simulator and both delivery pipelines have run, no band may have changed (dcegm_debug_check_guards);
  * racecheck: a data race that matters changes results from run to run or with the schedule — the same batch solved
    repeatedly, in batch order and costliest-first, chunked and whole, must give identical bytes every time.
(Out-of-bounds accesses and races that change a stored number, a length or a status word are also what every bitwise
comparison with the CPU oracle in the other GPU tests would trip over.)"""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def c3(n, na=500):
    from dcegm_jl_b200 import Param

    rng = np.random.default_rng(25032020)
    lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), size=4096)
    sig = rng.uniform(0, 0.6, size=4096)
    out = []
    for i in range(n):
        d = dict(na=na, beta=0.95, R=1.05, sigma=float(sig[i]), nsims=3000)
        d["lambda"] = float(lam[i])
        out.append(Param(par=d))
    return out


def guards(s):
    nb, bad = C.c_int(), C.c_int()
    assert s.lib.dcegm_debug_check_guards(s.h, C.byref(nb), C.byref(bad)) == 0
    return nb.value, bad.value


def test_no_kernel_writes_past_its_buffers():
    from dcegm_jl_b200 import Param, _abi
    from dcegm_jl_b200.solver import Solver

    with Solver(0) as s:
        ps = c3(320)
        s.solve(_abi.MODEL_F, ps, keep_dirty=True, fetch=False)           # resident retirement solver
        s.simulate(panel=True)
        s.solve_compact(_abi.MODEL_F, ps)                                 # one-launch pipeline, compact records
        s.solve_packed(_abi.MODEL_F, ps)                                  # ... (x, y) pairs
        s.solve(_abi.MODEL_F, c3(5, na=2400), fetch=False)                # fine-grid retirement solver
        s.fetch_compact()
        s.solve(_abi.MODEL_F, c3(3, na=3000), fetch=False)                # ... with the in-place simulator
        s.simulate()
        pg = [Param(par=dict(na=100, nsims=3000, rho=r, nu=nu)) for r in (0.0, 0.5, 0.95) for nu in (0.0, 3.0)]
        s.solve(_abi.MODEL_G, pg, keep_dirty=True, fetch=False)           # resident general-model solver
        s.simulate(panel=True)
        s.solve_compact(_abi.MODEL_G, pg * 60)                            # 360 general models through the pipeline
        s.solve(_abi.MODEL_G, pg, fine_grid=True, fetch=False)            # fine-grid general-model solver
        s.solve(_abi.MODEL_G, [Param(par=dict(na=1500, nT=8, nsims=2000))], fetch=False)
        s.simulate()
        for model in (_abi.MODEL_MINIMAL, _abi.MODEL_MINIMAL_BORROW, _abi.MODEL_MINIMAL_BEQUEST):
            s.solve(model, [Param(par=dict(na=100, a_low=-1.0 if model == _abi.MODEL_MINIMAL_BORROW else 0.0))] * 3, fetch=False)
            s.fetch_compact()
        nb, bad = guards(s)
        assert nb >= 20 and bad == 0, (nb, bad)


def test_results_do_not_depend_on_schedule_or_repetition():
    from dcegm_jl_b200 import _abi
    from dcegm_jl_b200.solver import Solver

    ps = c3(700)
    ref = None
    old = os.environ.get("DCEGM_LPT")
    try:
        for lpt in ("1", "0", "1"):
            os.environ["DCEGM_LPT"] = lpt  # read at upload: costliest-first or batch order
            with Solver(0) as s:
                for _ in range(2):
                    sol = s.solve(_abi.MODEL_F, ps)
                    summ = s.summaries()
                    blob = (sol.lines.tobytes(), sol.lens.tobytes(), sol.meta.tobytes(), sol.removed.tobytes(), sol.isects.tobytes(),
                            sol.vbound.tobytes(), sol.status.tobytes(), summ.tobytes())
                    if ref is None:
                        ref = blob
                    assert blob == ref
                cl = s.solve_compact(_abi.MODEL_F, ps)  # one launch, per-chunk order, encoder running beside the solver
                sol2 = s.fetch()
                assert (sol2.lines.tobytes(), sol2.lens.tobytes(), sol2.status.tobytes()) == (ref[0], ref[1], ref[6])
                os.environ["DCEGM_ONE_LAUNCH"] = "0"
                cl2 = s.solve_compact(_abi.MODEL_F, ps)  # one launch per chunk
                os.environ.pop("DCEGM_ONE_LAUNCH")
                # other chunk boundaries, so other models carry the shared retiree records: same lines all the same
                assert cl2.expand()[0].tobytes() == cl.expand()[0].tobytes()
                assert np.array_equal(cl2.aux, cl.aux) and np.array_equal(cl2.lens, cl.lens)
    finally:
        os.environ.pop("DCEGM_ONE_LAUNCH", None)
        if old is None:
            os.environ.pop("DCEGM_LPT", None)
        else:
            os.environ["DCEGM_LPT"] = old
