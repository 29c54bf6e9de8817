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

from bench_configs import c3_params, run_c2, run_c4, run_c5, status_counts  # noqa: E402

NA, NT, NY = 500, 25, 5
BATCH = 4096
# 2 lines x 2 choices x 24 periods x 501 points x 16 B + terminal 128 B + param 248 B
ALGO_BYTES_PER_MODEL = 2 * 2 * 24 * 501 * 16 + 128 + 248
WORKLOAD = "C3: retirement model (FModel) batch, na=500, nT=25, ny=5, log10(lambda)~U(log10 2e-6, log10 0.2), sigma~U(0,0.6)"
KERNEL_SOURCES = ["solver_f.cu", "device_common.cuh", "egm_node.cuh", "egm_point.inc", "envelope.cuh", "dc_math.h", "Makefile"]
COUNTERS_FILE = os.path.join(ROOT, "profiles", "solve_f_counters.json")


def kernel_source_hash():
    """sha256 over the sources solve_f_kernel is built from: ties the ncu counters under profiles/ to a build
    (the GPU box has no .git, so a content hash stands in for the commit)"""
    import hashlib

    h = hashlib.sha256()
    for f in KERNEL_SOURCES:
        with open(os.path.join(ROOT, "dcegm.jl_b200", "csrc", f), "rb") as fh:
            data = fh.read()
        if f == "Makefile":  # only what decides how a kernel is compiled, not the list of files linked into the library
            data = b"".join(ln for ln in data.splitlines(True) if ln.split(b":=")[0].split(b"?=")[0].strip() in (b"NVCC", b"ARCH", b"NVFLAGS"))
        h.update(data)
    return h.hexdigest()[:16]


def ncu_counters():
    """per-model ncu counters of solve_f_kernel (DRAM bytes, FP64-pipe warp instructions) — only if they were
    captured on exactly the kernel sources of this build (tools/ncu_counters_json.py writes the file)"""
    try:
        with open(COUNTERS_FILE) as f:
            c = json.load(f)
    except Exception:
        return None, "no profiles/solve_f_counters.json"
    now = kernel_source_hash()
    if c.get("source_sha16") != now:
        return None, f"profiles/solve_f_counters.json was captured on sources {c.get('source_sha16')}, this build is {now}: stale, not printed"
    return c, f"ncu capture {c.get('capture')} on sources {now} ({c.get('models')} models)"


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
    """The CPU restatement of the reference path (oracle/), `threads` host threads.  Returns models/s."""
    from dcegm_jl_b200 import _abi
    from oracle import oracle

    arr = c3_params(nmodels, offset=seed_offset)
    oracle.lib()
    t0 = time.perf_counter()
    sol = oracle.solve(_abi.MODEL_F, arr, nthreads=threads)
    dt = time.perf_counter() - t0
    ok, _ = status_counts(sol.status)  # models on which the reference algorithm would throw do not count
    return ok / dt, dt


def parity_in_run(solver, arr, threads, chunk=256):
    """Checker leg (outside every timed region): the benchmarked batch, resident on the GPU from the timed solves,
    against the CPU oracle in portable-math mode — status words, line lengths, envelope metadata, removed-point
    bitmasks and every stored point of every model, bit for bit."""
    from dcegm_jl_b200 import _abi
    from dcegm_jl_b200.param import params_array
    from oracle import oracle

    n = len(arr)
    out = {"models": n, "oracle_mode": "portable (dc_math.h arithmetic)", "status_equal": True, "lens_equal": True,
           "meta_equal": True, "removed_equal": True, "points_compared": 0, "points_differ": 0}
    t0 = time.perf_counter()
    for c0 in range(0, n, chunk):
        cnt = min(chunk, n - c0)
        ref = oracle.solve(_abi.MODEL_F, params_array([arr[i] for i in range(c0, c0 + cnt)]), nthreads=threads, portable=True)
        gpu = solver.fetch(c0, cnt)
        out["status_equal"] &= bool(np.array_equal(gpu.status, ref.status))
        out["lens_equal"] &= bool(np.array_equal(gpu.lens, ref.lens))
        out["meta_equal"] &= bool(np.array_equal(gpu.meta, ref.meta))
        out["removed_equal"] &= bool(np.array_equal(gpu.removed, ref.removed))
        k = np.arange(ref.lay.cap)
        m = np.broadcast_to((k[None, None, None, None, None, :] < ref.lens[..., None])[..., None], ref.lines.shape)
        x, y = gpu.lines[m].view(np.int64), ref.lines[m].view(np.int64)
        out["points_compared"] += int(x.size // 2)
        out["points_differ"] += int((x != y).reshape(-1, 2).any(axis=1).sum())
    out["seconds"] = round(time.perf_counter() - t0, 1)
    out["all_equal"] = bool(out["status_equal"] and out["lens_equal"] and out["meta_equal"] and out["removed_equal"]
                            and out["points_differ"] == 0)
    return out


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
    v1, dt1 = cpu_reference_run(48, 1)
    out = {
        "impl": "reference", "metric": "models solved/sec (500pts x 25T retirement model)", "value": v,
        "unit": "models/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "models_per_step": sample},
        "cpu_baseline": {"value": v, "unit": "models/s", "cores": threads, "kind": "port",
                         "sample": f"{sample} models of the C3 batch per step, C++ restatement of DCEGM.jl "
                                   f"(julia unavailable on this box), one model per task on {threads} threads",
                         "one_thread": {"value": v1, "unit": "models/s", "cores": 1,
                                        "sample": f"first 48 models of the batch, {dt1:.1f} s; the published Julia figure is "
                                                  "7.94 models/s on one laptop core (README.md:118-128)"}},
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
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline and parity_in_run legs")
    ap.add_argument("--no-configs", action="store_true", help="skip the C2 / C4 / C5 lines")
    ap.add_argument("--na", type=int, default=NA, help="savings grid points (experiments only; the metric is quoted at 500)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    # stdout carries exactly one JSON line: libraries that print there while the job runs (NCCL's version banner on
    # some boxes) go to stderr instead
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)

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
    n_ok, st_bits = status_counts(st)
    # models with a fatal status bit (the reference algorithm itself fails there) are not counted as solved
    t = torch.tensor([float(n_ok)], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    solved = float(t.item())
    value = solved * args.steps / (tot_ms_max * 1e-3)

    # ---- end to end through the C ABI with host buffers ("e2e") ------------------------------
    # host Param array -> dcegm_solve_f_compact (host model setup + H2D + kernels) -> every stored polyline delivered
    # into pinned host memory in the compact line format (include/dcegm.h: ordinates + abscissa runs, lossless),
    # plus offsets / lens / run counts / status.
    nslots = nper * lay.nT * lay.nstate * 4
    woff = np.zeros(nslots + 1, np.int64)
    lens_h = np.zeros(nslots, np.int32)
    aux_h = np.zeros(nslots, np.int32)
    stat_h = np.zeros(nper, np.int32)
    nw = C.c_int64()
    lp, ip = C.POINTER(C.c_int64), C.POINTER(C.c_int32)
    wp, lnp, axp = woff.ctypes.data_as(lp), lens_h.ctypes.data_as(ip), aux_h.ctypes.data_as(ip)
    assert lib.dcegm_fetch_compact(h, 0, nper, None, 0, wp, lnp, axp, C.byref(nw)) == 0, lib.dcegm_last_error(h)
    cap_words = int(nw.value * 1.05) + 4096
    npts_total = int(lens_h.sum())
    pinned = C.c_void_p()
    assert lib.dcegm_host_alloc(C.byref(pinned), max(cap_words * 8, npts_total * 16 + 4096)) == 0
    pin_d = C.cast(pinned, C.POINTER(C.c_double))
    h2d = C.sizeof(arr) + 8 * (lay.nT - 1) * 2 * lay.ny * nper + 8 * lay.ny * nper + 8 * lay.na + 200 * nper

    def e2e_step():
        # one public call: host setup + H2D + chunked solve; encoding and D2H overlapped with the kernels
        rc = lib.dcegm_solve_f_compact(h, arr, nper, 0, pin_d, cap_words, wp, lnp, axp, C.byref(nw))
        assert rc == 0, lib.dcegm_last_error(h)
        lib.dcegm_status(h, stat_h.ctypes.data_as(ip))
        return nw.value * 8 + (nslots + 1) * 8 + nslots * 8 + nper * 4

    def e2e_xy_step():
        # the same with every point delivered as (x, y) pairs (Vector{Point} layout, 16 B per point)
        npts = C.c_int64()
        rc = lib.dcegm_solve_f_packed(h, arr, nper, 0, pin_d, npts_total + 256, wp, C.byref(npts))
        assert rc == 0, lib.dcegm_last_error(h)
        lib.dcegm_status(h, stat_h.ctypes.data_as(ip))
        return npts.value * 16 + (nslots + 1) * 8 + nper * 4

    def timed_e2e(fn, steps):
        for _ in range(2):
            fn()
        if dist:
            dist.barrier()
        t0 = time.perf_counter()
        nbytes, step_ms = 0, []
        for _ in range(steps):
            t1 = time.perf_counter()
            nbytes = fn()
            step_ms.append((time.perf_counter() - t1) * 1e3)
        secs = time.perf_counter() - t0
        if os.environ.get("DCEGM_TRACE"):
            print("e2e step wall ms:", [round(x, 2) for x in step_ms], file=sys.stderr)
        tt = torch.tensor([secs], dtype=torch.float64, device="cuda")
        if dist:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return solved * steps / float(tt.item()), nbytes, nbytes * steps / secs / 1e9

    e2e_xy_value, d2h_xy, _ = timed_e2e(e2e_xy_step, min(args.steps, 3))
    e2e_value, d2h, d2h_gbs = timed_e2e(e2e_step, args.steps)
    e2e_launches = int(solver.timing().n_launches)  # kernels of one end-to-end step: solver + encoder passes per chunk
    # slots delivered as a reference to an identical line of their chunk's first model (no record of their own)
    idx_ = np.flatnonzero(lens_h > 0)
    _, first_ = np.unique(woff[:-1][idx_], return_index=True)  # the first slot at an offset owns the record
    own_ = np.zeros(len(idx_), bool)
    own_[first_] = True
    shared_slots = int((~own_).sum())
    shared_points = int(lens_h[idx_[~own_]].sum())
    clocks = sampler.stop() if rank == 0 else None  # sampled over both timed regions (device-resident and e2e)
    gbs_all = None
    if dist:  # per-rank device-to-host rates (the host side limits e2e on a multi-GPU box)
        g = torch.zeros(world, dtype=torch.float64, device="cuda")
        g[rank] = d2h_gbs
        dist.all_reduce(g, op=dist.ReduceOp.SUM)
        gbs_all = [round(float(x), 2) for x in g.tolist()]

    # expansion of the delivered records into (x, y) points on the host threads (not part of e2e: the host mirror
    # decodes lines lazily, per Envelope); checked here against the resident solution's packed points
    expand = None
    if rank == 0:
        xy = np.empty((npts_total, 2))
        off = np.zeros(nslots + 1, np.int64)
        npts = C.c_int64()
        threads = os.cpu_count() or 1
        t_exp = []
        for _ in range(2):  # the first pass also faults in the 3.15 GB destination; the second is the decoder alone
            t0 = time.perf_counter()
            rc = lib.dcegm_expand_compact(h, 0, nper, pin_d, wp, lnp, axp, xy.ctypes.data_as(C.POINTER(C.c_double)), npts_total,
                                          off.ctypes.data_as(lp), C.byref(npts), threads)
            t_exp.append(time.perf_counter() - t0)
            assert rc == 0, lib.dcegm_last_error(h)
        ref_xy, ref_off = solver.fetch_packed()
        expand = {"ms": t_exp[1] * 1e3, "ms_first_touch": t_exp[0] * 1e3, "threads": threads, "points": int(npts.value),
                  "equals_packed_points_bitwise": bool(xy.tobytes() == ref_xy.tobytes() and np.array_equal(off, ref_off))}
        del xy, ref_xy
    lib.dcegm_host_free(pinned)

    configs = None
    if not args.no_configs:
        configs = {}
        if world == 1:
            configs["c2"] = run_c2(solver)
            configs["c4"] = run_c4(solver)
        configs["c5"] = run_c5(solver, local, world, rank, dist)
        solver.upload(_abi.MODEL_F, arr)  # the benchmarked batch back into residence for the parity leg
        solver.run()

    if rank == 0:
        peak, peak_src = measured_peak()
        ms_kernel = kern_ms / args.steps
        # FP64 issue rate of this GPU, measured now (independent DFMA chains at full occupancy)
        dfma, dfma_ms = C.c_double(), C.c_double()
        lib.dcegm_debug_dfma_rate(local, C.byref(dfma), C.byref(dfma_ms))
        fp64_peak = dfma.value / 32 / 1e9
        achieved = ALGO_BYTES_PER_MODEL * nper / (ms_kernel * 1e-3) / 1e9  # bytes of one launch (rank 0) / its duration
        ctr, ctr_src = ncu_counters()
        traffic = fp64_achieved = None
        if ctr:
            traffic = (ctr["dram_bytes_read"] + ctr["dram_bytes_write"]) / ctr["models"] * nper
            fp64_achieved = ctr["fp64_warp_inst"] / ctr["models"] * nper / (ms_kernel * 1e-3) / 1e9
        out = {
            "metric": "models solved/sec (500pts x 25T retirement model)", "value": value, "unit": "models/s",
            "n_gpus": world, "steps": args.steps, "warmup": W, "ms_per_step": tot_ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": value / 7.94,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "models_per_gpu": nper, "global_batch": nper * world,
                       "l2": "256 MiB flush between timed steps; each step also writes 3.2 GB of lines (>> 126 MB L2)",
                       "vs_baseline_ref": "README.md:118-128: 125.947 ms median per solve = 7.94 models/s, 1 laptop core, Julia",
                       "models_counted": int(solved), "flagged_models_rank0": nper - n_ok, "status_bits_rank0": st_bits,
                       "flagged_note": "models with a fatal status bit are not counted as solved; at C3 all are DCEGM_ST_DUP_KNOT: the "
                                       "reference's upper_env pushes an intersection found at a shared turning point twice "
                                       "(envelope.jl:199, :212, :248-252; DESIGN.md §2)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": ctr_src + ": dram__bytes_read.sum + dram__bytes_write.sum per launch, scaled to this batch",
                         "peak_source": peak_src,
                         "note": "dominant kernel solve_f_kernel is FP64-pipe bound by construction (SURVEY.md §8d): "
                                 "~70 FP64 instr per algorithmic byte; HBM fraction reported as the metric asks"},
            "roofline_fp64": {"bound": "fp64 pipe", "achieved": fp64_achieved, "peak": fp64_peak, "unit": "G warp-inst/s",
                              "frac": fp64_achieved / fp64_peak if fp64_achieved and fp64_peak > 0 else None,
                              "peak_tflops": 2 * dfma.value / 1e12,
                              "source": "achieved = FP64-pipe warp instructions per model (sm__inst_executed_pipe_fp64.sum; " + ctr_src +
                                        ") x models / kernel time; peak = dcegm_debug_dfma_rate measured in this run"},
            "e2e": {"value": e2e_value, "unit": "models/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "d2h_gbs": d2h_gbs, "d2h_gbs_per_rank": gbs_all, "gpu_launches_per_step": e2e_launches,
                    "format": "compact lines (include/dcegm.h): ordinates + abscissa runs, lossless; a line bit-identical to the same "
                              "slot of its chunk's first model travels once; expanded lazily per line by the host mirror",
                    "lines_delivered_by_reference": {"slots": shared_slots, "of_slots": int((lens_h > 0).sum()), "points": shared_points,
                                                     "note": "the retiree's lines do not depend on (lambda, sigma), the two swept "
                                                             "parameters of C3; DCEGM_COMPACT_ALIAS=0 ships every line on its own"},
                    "expand_all_on_host": expand,
                    "full_xy": {"value": e2e_xy_value, "unit": "models/s", "d2h_bytes_per_step": int(d2h_xy),
                                "note": "the same call delivering every point as an (x, y) pair (dcegm_solve_f_packed)"},
                    "note": "the solver kernel with the encoder running beside it, plus upload and the last chunk's encode + copy"},
            "gpu_launches": args.steps * 1,
            "clocks": clocks,
        }
        if configs is not None:
            out["configs"] = configs
        if not args.no_cpu:
            threads = os.cpu_count() or 1
            sample = min(nper, BATCH, 512 * threads)  # ~10-20 s of CPU work
            v, dt = cpu_reference_run(sample, threads)
            v1, dt1 = cpu_reference_run(48, 1)
            out["cpu_baseline"] = {"value": v, "unit": "models/s", "cores": threads, "kind": "port",
                                   "sample": f"first {sample} models of the same C3 batch, C++ restatement of DCEGM.jl "
                                             f"(oracle/; julia unavailable on this box), {threads} host threads, {dt:.1f} s",
                                   "one_thread": {"value": v1, "unit": "models/s", "cores": 1,
                                                  "sample": f"first 48 models, {dt1:.1f} s; published Julia figure: 7.94 models/s "
                                                            "on one laptop core (README.md:118-128)"}}
            if args.na == NA:
                out["parity_in_run"] = parity_in_run(solver, arr, threads)
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(out), flush=True)
        os.dup2(2, 1)
    solver.close()
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
