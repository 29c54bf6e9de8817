"""Synthetic draws of the BASELINE.json configurations and device-timed runs of the ones
bench.py reports beside its headline workload (C2, C4, C5).  Imported by bench.py and tools/."""
import ctypes as C
import json
import os
import time

import numpy as np

SEED = 25032020  # the reference's own rseed (param.json:137-141)


def c3_params(n, seed=SEED, offset=0, na=500):
    """C3: n retirement models, na x 25 periods, ny 5, log10(lambda) ~ U(log10 2e-6, log10 0.2), sigma ~ U(0, 0.6)"""
    from dcegm_jl_b200 import Param
    from dcegm_jl_b200.param import params_array

    rng = np.random.default_rng(seed)
    lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), size=offset + n)
    sig = rng.uniform(0, 0.6, size=offset + n)
    base = Param(par=dict(na=na, beta=0.95, R=1.05, nT=25, ny=5)).to_c()
    arr = params_array([base] * n)
    for i in range(n):
        arr[i].lambda_ = float(lam[offset + i])
        arr[i].sigma = float(sig[offset + i])
    return arr


def c2_params(n=4096, seed=SEED):
    """C2: minimal consumption-savings EGM, beta ~ U(0.5, 1), R ~ U(1, 1.5) (slider ranges interact.jl:5-6)"""
    from dcegm_jl_b200 import Param
    from dcegm_jl_b200.param import params_array

    rng = np.random.default_rng(seed)
    base = Param(par=dict(sigma=0.35, na=100, ny=5, nT=25, a_high=50.0, cfloor=1e-3)).to_c()
    arr = params_array([base] * n)
    for i, (b, r) in enumerate(zip(rng.uniform(0.5, 1.0, n), rng.uniform(1.0, 1.5, n))):
        arr[i].beta, arr[i].R = float(b), float(r)
    return arr


def c4_params(n=256, seed=SEED):
    """C4: fine-grid retirement model, na 20000, nT 50, ny 32, retage 51, (lambda, sigma) as C3"""
    from dcegm_jl_b200 import Param
    from dcegm_jl_b200.param import params_array

    rng = np.random.default_rng(seed)
    base = Param(par=dict(na=20000, nT=50, ny=32, retage=51, beta=0.95, R=1.05)).to_c()
    arr = params_array([base] * n)
    lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), n)
    sig = rng.uniform(0, 0.6, n)
    for i in range(n):
        arr[i].lambda_, arr[i].sigma = float(lam[i]), float(sig[i])
    return arr


def c5_params(n=8192, nsims=100000, seed=SEED, rank=0):
    """C5 shard: general models (GModel, na 100, ny 5, nT 25), ranges of interact.jl:106-116; every rank draws its own"""
    from dcegm_jl_b200 import Param
    from dcegm_jl_b200.param import params_array

    rng = np.random.default_rng(seed + rank)
    base = Param(par=dict(na=100, nsims=nsims)).to_c()
    arr = params_array([base] * n)
    for i in range(n):
        a = arr[i]
        a.beta, a.R, a.alpha = float(rng.uniform(0.5, 1)), float(rng.uniform(1, 1.5)), float(rng.uniform(0, 0.5))
        a.sigma, a.rho, a.pension = float(rng.uniform(0, 0.35)), float(rng.uniform(0, 0.95)), float(rng.uniform(0, 1))
        a.lambda_ = float(10 ** rng.uniform(np.log10(2e-7), 0))
    return arr


def status_counts(st):
    """(models without a fatal status bit, {bit name: models}) — only the former count as solved"""
    from dcegm_jl_b200 import _abi

    st = np.asarray(st).astype(np.int64) & 0xFFFFFFFF
    names = {b: k for k, b in vars(_abi).items() if k.startswith("ST_") and isinstance(b, int) and b and (b & (b - 1)) == 0}
    bits = {names[1 << b]: int(((st >> b) & 1).sum()) for b in range(16) if ((st >> b) & 1).any()}
    return int(((st & _abi.ST_FATAL_MASK) == 0).sum()), bits


def run_c2(solver, reps=5):
    from dcegm_jl_b200 import _abi

    arr = c2_params()
    solver.upload(_abi.MODEL_MINIMAL, arr)
    solver.run()
    ms = []
    for _ in range(reps):
        solver.run()
        ms.append(solver.timing().kernel_ms)
    ok, bits = status_counts(solver.status())
    t = float(np.median(ms))
    return {"workload": "C2: minimal EGM x4096 (na=100, ny=5, nT=25)", "models": len(arr), "solved": ok, "status_bits": bits,
            "kernel_ms": t, "reps": reps, "models_per_s": ok / (t * 1e-3), "launches_per_rep": solver.timing().n_launches}


def run_c4(solver, n=256, reps=3):
    from dcegm_jl_b200 import _abi

    arr = c4_params(n)
    solver.upload(_abi.MODEL_F, arr)
    solver.run()
    ms = []
    for _ in range(reps):
        solver.run()
        ms.append(solver.timing().kernel_ms)
    ok, bits = status_counts(solver.status())
    t = float(np.median(ms))
    algo = 2 * 2 * 49 * 20001 * 16 * n
    out = {"workload": f"C4: fine-grid retirement model x{n} (na=20000, nT=50, ny=32)", "models": n, "solved": ok,
           "status_bits": bits, "kernel_ms": t, "reps": reps, "models_per_s": ok / (t * 1e-3),
           "launches_per_rep": solver.timing().n_launches, "algorithmic_gbs": algo / (t * 1e-3) / 1e9}
    # end to end through dcegm_solve_f_compact: host Param array in, every stored polyline in pinned host memory out
    lib, h = solver.lib, solver.h
    lay = solver.layout()
    nslots = n * lay.nT * lay.nstate * 4
    woff = np.zeros(nslots + 1, np.int64)
    lens = np.zeros(nslots, np.int32)
    aux = np.zeros(nslots, np.int32)
    nw = C.c_int64()
    lp, ip = C.POINTER(C.c_int64), C.POINTER(C.c_int32)
    args = (woff.ctypes.data_as(lp), lens.ctypes.data_as(ip), aux.ctypes.data_as(ip), C.byref(nw))
    if lib.dcegm_fetch_compact(h, 0, n, None, 0, *args) == 0:
        cap_words = int(nw.value * 1.05) + 4096
        pinned = C.c_void_p()
        if lib.dcegm_host_alloc(C.byref(pinned), cap_words * 8) == 0:
            ts = []
            for _ in range(3):
                t0 = time.perf_counter()
                rc = lib.dcegm_solve_f_compact(h, arr, n, 0, C.cast(pinned, C.POINTER(C.c_double)), cap_words, *args)
                ts.append(time.perf_counter() - t0)
                if rc != 0:
                    break
            lib.dcegm_host_free(pinned)
            if rc == 0:
                out["e2e_ms"] = min(ts[1:]) * 1e3
                out["e2e_d2h_bytes"] = int(nw.value * 8 + woff.nbytes + lens.nbytes + aux.nbytes)
                out["e2e_solved_models_per_s"] = ok / min(ts[1:])
                out["e2e_xy_bytes"] = int(lens.astype(np.int64).sum() * 16)
    return out


def run_c5(solver, local, world, rank, dist, per=8192, nsims=100000, reps=3):
    """C5 as one timed job per repetition: solve the rank's shard of general models, simulate nsims agents per model,
    ONE all-gather of the per-model records (summary + per-period moments); timed with CUDA events on the torch
    stream around the whole job (the library's stream work is awaited inside), max over ranks."""
    import torch

    from dcegm_jl_b200 import _abi, sweep

    arr = c5_params(per, nsims, rank=rank)
    nT = 25
    rd = sweep.record_dtype(nT)
    rec = rd.itemsize
    dev = torch.device("cuda", local)
    send = torch.empty(rec * per, dtype=torch.uint8, device=dev)
    n = per * world

    def job():
        solver.solve(_abi.MODEL_G, arr, fetch=False)
        kms = solver.timing().kernel_ms
        solver.simulate()
        solver.sweep_records_to_device(send.data_ptr(), rec * per, nT)
        out = sweep.gather_records(send, n, rec, dist if world > 1 else None)
        torch.cuda.synchronize()
        return out, kms

    out, kms = job()  # warm-up: allocations, communicator
    secs = []
    for _ in range(reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out, kms = job()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        secs.append(float(dt.item()))
    recs = np.frombuffer(out.cpu().numpy().tobytes(), dtype=rd)
    ok, bits = status_counts(recs["summary"]["status"])
    t = float(np.median(secs))
    return {"workload": f"C5: {n} general models (na=100, ny=5, nT=25) + {nsims} simulated agents each on {world} GPU(s), "
                        "one all-gather of summaries + moments", "models": n, "solved": ok, "status_bits": bits,
            "seconds": t, "reps": reps, "models_per_s": ok / t, "agent_periods_per_s": n * nsims * nT / t,
            "solve_kernel_ms_rank0": kms, "allgather_bytes": int(out.numel()), "collectives_per_job": 1 if world > 1 else 0,
            "masked_moment_models": int(np.isnan(recs["moments"][:, 0, 2]).sum())}
