#!/usr/bin/env python
"""bench.py — models solved / s on the retirement-model batch (BASELINE.json configs[2], "C3":
4096 (lambda, sigma) draws, 500 savings points, 25 periods) on N B200s.

A step = one backward-induction solve of the rank's whole batch (EGM step + upper envelope for
every period and choice).  `value` is timed with CUDA events on the library's stream with all inputs
resident in HBM; `e2e` goes through the public C-ABI call with host buffers (host setup, H2D, solve,
D2H of every stored polyline into pinned host memory).  See DESIGN.md §Measurement.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

NA, NT, NY = 500, 25, 5
BATCH = 4096
# 2 lines x 2 choices x 24 periods x 501 points x 16 B + terminal 128 B + param 248 B
ALGO_BYTES_PER_MODEL = 2 * 2 * 24 * 501 * 16 + 128 + 248
# dram__bytes_read.sum + dram__bytes_write.sum of solve_f_kernel per model, from the `ncu --set full` capture
# summarised in profiles/r1g_ncu_summary.md (11.53 MB + 885.39 MB for 1184 models)
NCU_DRAM_BYTES_PER_MODEL = (11.529984e6 + 885.387008e6) / 1184
# sm__inst_executed_pipe_fp64.sum of solve_f_kernel per model (warp instructions), same capture series
# (profiles/r1g_solve_f_metrics.csv: 978 056 424 for 1184 models)
NCU_FP64_WARP_INST_PER_MODEL = 978056424 / 1184
WORKLOAD = "C3: retirement model (FModel) batch, na=500, nT=25, ny=5, log10(lambda)~U(log10 2e-6, log10 0.2), sigma~U(0,0.6)"


def c3_params(n, seed=25032020, offset=0, na=None):
    from dcegm_jl_b200 import Param
    from dcegm_jl_b200.param import params_array

    rng = np.random.default_rng(seed)
    lam = 10 ** rng.uniform(np.log10(2e-6), np.log10(0.2), size=offset + n)
    sig = rng.uniform(0, 0.6, size=offset + n)
    base = Param(par=dict(na=na or NA, beta=0.95, R=1.05, nT=NT, ny=NY)).to_c()
    arr = params_array([base] * n)
    for i in range(n):
        arr[i].lambda_ = float(lam[offset + i])
        arr[i].sigma = float(sig[offset + i])
    return arr


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)"""

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_reference_run(nmodels, threads, seed_offset=0):
    """The CPU restatement of the reference path (oracle/), all host threads.  Returns models/s."""
    from dcegm_jl_b200 import _abi
    from oracle import oracle

    arr = c3_params(nmodels, offset=seed_offset)
    oracle.lib()
    t0 = time.perf_counter()
    sol = oracle.solve(_abi.MODEL_F, arr, nthreads=threads)
    dt = time.perf_counter() - t0
    ok = int((sol.status == 0).sum())  # models on which the reference algorithm would throw do not count
    return ok / dt, dt


def run_reference(args):
    """--impl reference: the reference's own algorithm on the host cores.  DCEGM.jl itself cannot run
    (no julia on this box): the timed code is the C++ restatement under oracle/ (kind "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = min(BATCH, 256 * threads)  # a few seconds of CPU work per step
    for _ in range(args.warmup):
        cpu_reference_run(min(sample, 4 * threads), threads)
    tot_t, tot_n = 0.0, 0
    for k in range(args.steps):
        v_, dt = cpu_reference_run(sample, threads, seed_offset=(k * sample) % BATCH)
        tot_t += dt
        tot_n += v_ * dt
    v = tot_n / tot_t
    out = {
        "impl": "reference", "metric": "models solved/sec (500pts x 25T retirement model)", "value": v,
        "unit": "models/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "models_per_step": sample},
        "cpu_baseline": {"value": v, "unit": "models/s", "cores": threads, "kind": "port",
                         "sample": f"{sample} models of the C3 batch per step, C++ restatement of DCEGM.jl "
                                   f"(julia unavailable on this box), one model per task on {threads} threads"},
        "e2e": {"value": v, "unit": "models/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--batch", type=int, default=BATCH, help="models per GPU")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--na", type=int, default=NA, help="savings grid points (experiments only; the metric is quoted at 500)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch

    from dcegm_jl_b200 import _abi
    from dcegm_jl_b200.solver import Solver, SUMMARY_DTYPE

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libdcegm_cuda has no CPU path")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W = max(args.warmup, 3)
    nper = args.batch
    arr = c3_params(nper, offset=rank * nper, na=args.na)  # contiguous shard of the global draw
    solver = Solver(local)
    lib, h = solver.lib, solver.h

    # ---- device-resident timing ("value") ----------------------------------------------------
    solver.upload(_abi.MODEL_F, arr)
    lay = solver.layout()
    sbytes = SUMMARY_DTYPE.itemsize * nper
    send = torch.empty(sbytes, dtype=torch.uint8, device="cuda")
    recv = torch.empty(sbytes * world, dtype=torch.uint8, device="cuda") if world > 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def step():
        flush.zero_()  # L2 flush between timed iterations (outside the timed events)
        torch.cuda.synchronize()
        solver.run()  # CUDA events on the library's stream bracket the solve
        ms = solver.timing().kernel_ms
        if world > 1:  # the one collective of a sweep: allgather of the per-model summaries
            solver.summaries_to_device(send.data_ptr(), sbytes)
            e0.record()
            dist.all_gather_into_tensor(recv, send)
            e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        return ms

    for _ in range(W):
        step()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    tot_ms, kern_ms = 0.0, 0.0
    for _ in range(args.steps):
        tot_ms += step()
        kern_ms += solver.timing().kernel_ms
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    t = torch.tensor([tot_ms], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot_ms_max = float(t.item())
    st = solver.status()
    n_flagged = int((st != 0).sum())
    # models flagged by status (the reference algorithm itself would throw there) are not
    # counted as solved
    t = torch.tensor([float(nper - n_flagged)], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    solved = float(t.item())
    value = solved * args.steps / (tot_ms_max * 1e-3)

    # ---- end to end through the C ABI with host buffers ("e2e") ------------------------------
    # host Param array -> dcegm_solve_f (host model setup + H2D + kernels) -> every stored polyline
    # copied densely packed into pinned host memory, plus lens/meta/status.
    nslots = nper * lay.nT * lay.nstate * 4
    offsets = np.zeros(nslots + 1, np.int64)
    npts = C.c_int64()
    op = offsets.ctypes.data_as(C.POINTER(C.c_int64))
    lib.dcegm_fetch_packed(h, 0, nper, None, 0, op, C.byref(npts))
    cap_pts = int(npts.value * 1.05) + 1024
    pinned = C.c_void_p()
    assert lib.dcegm_host_alloc(C.byref(pinned), cap_pts * 16) == 0
    lens_h = np.zeros(nslots, np.int32)
    stat_h = np.zeros(nper, np.int32)
    h2d = C.sizeof(arr) + 8 * (lay.nT - 1) * 2 * lay.ny * nper + 8 * lay.ny * nper + 8 * lay.na + 200 * nper

    def e2e_step():
        # one public call: host setup + H2D + chunked solve, packing and D2H overlapped with the kernels
        rc = lib.dcegm_solve_f_packed(h, arr, nper, 0, C.cast(pinned, C.POINTER(C.c_double)), cap_pts, op, C.byref(npts))
        assert rc == 0, lib.dcegm_last_error(h)
        lib.dcegm_status(h, stat_h.ctypes.data_as(C.POINTER(C.c_int32)))
        return npts.value * 16 + nslots * 4 + nper * 4

    for _ in range(2):
        e2e_step()
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    d2h = 0
    step_ms = []
    for _ in range(args.steps):
        t1 = time.perf_counter()
        d2h = e2e_step()
        step_ms.append((time.perf_counter() - t1) * 1e3)
    e2e_s = time.perf_counter() - t0
    if os.environ.get("DCEGM_TRACE"):
        print("e2e step wall ms:", [round(x, 2) for x in step_ms], file=sys.stderr)
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = solved * args.steps / float(t.item())
    clocks = sampler.stop() if rank == 0 else None  # sampled over both timed regions (device-resident and e2e)
    lib.dcegm_host_free(pinned)

    if rank == 0:
        peak, peak_src = measured_peak()
        ms_kernel = kern_ms / args.steps
        # FP64 issue rate of this GPU, measured now (independent DFMA chains at full occupancy)
        dfma, dfma_ms = C.c_double(), C.c_double()
        lib.dcegm_debug_dfma_rate(local, C.byref(dfma), C.byref(dfma_ms))
        fp64_achieved = NCU_FP64_WARP_INST_PER_MODEL * nper / (ms_kernel * 1e-3) / 1e9
        fp64_peak = dfma.value / 32 / 1e9
        achieved = ALGO_BYTES_PER_MODEL * nper / (ms_kernel * 1e-3) / 1e9  # bytes of one launch (rank 0) / its duration
        out = {
            "metric": "models solved/sec (500pts x 25T retirement model)", "value": value, "unit": "models/s",
            "n_gpus": world, "steps": args.steps, "warmup": W, "ms_per_step": tot_ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": value / 7.94,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "models_per_gpu": nper, "global_batch": nper * world,
                       "l2": "256 MiB flush between timed steps; each step also writes 3.2 GB of lines (>> 126 MB L2)",
                       "vs_baseline_ref": "README.md:118-128: 125.947 ms median per solve = 7.94 models/s, 1 laptop core, Julia",
                       "flagged_models_rank0": n_flagged, "models_counted": int(solved),
                       "flagged_note": "status != 0 (non-monotone line: Interpolations.jl would throw) -> not counted; the oracle flags the same models"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_DRAM_BYTES_PER_MODEL * nper, "traffic_source": "ncu dram__bytes_read.sum + dram__bytes_write.sum (profiles/r1g_ncu_summary.md, r1g_solve_f_metrics.csv), bytes per launch scaled to this batch",
                         "peak_source": peak_src,
                         "note": "dominant kernel solve_f_kernel is FP64-pipe bound by construction (SURVEY.md §8d): "
                                 "~70 FP64 instr per algorithmic byte; HBM fraction reported as the metric asks"},
            "roofline_fp64": {"bound": "fp64 pipe", "achieved": fp64_achieved, "peak": fp64_peak, "unit": "G warp-inst/s",
                              "frac": fp64_achieved / fp64_peak if fp64_peak > 0 else None,
                              "peak_tflops": 2 * dfma.value / 1e12,
                              "source": "achieved = FP64-pipe warp instructions per model (ncu sm__inst_executed_pipe_fp64.sum, "
                                        "profiles/r1g_solve_f_metrics.csv) x models / kernel time; peak = dcegm_debug_dfma_rate measured in this run"},
            "e2e": {"value": e2e_value, "unit": "models/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "d2h_gbs": d2h * args.steps / e2e_s / 1e9,
                    "note": "bound by the device-to-host transfer of every stored polyline (pinned D2H measures ~57 GB/s on these boxes)"},
            "gpu_launches": args.steps * 1,
            "clocks": clocks,
        }
        if not args.no_cpu:
            threads = os.cpu_count() or 1
            sample = min(BATCH, 512 * threads)  # ~10-20 s of CPU work
            v, dt = cpu_reference_run(sample, threads)
            out["cpu_baseline"] = {"value": v, "unit": "models/s", "cores": threads, "kind": "port",
                                   "sample": f"first {sample} models of the same C3 batch, C++ restatement of DCEGM.jl "
                                             f"(oracle/; julia unavailable on this box), {threads} host threads, {dt:.1f} s"}
        print(json.dumps(out))
    solver.close()
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
