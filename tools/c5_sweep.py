"""BASELINE.json configs[4] ("C5") as a timed multi-GPU job, run under torchrun (one rank per GPU): every rank
solves its shard of state-dependent general models (GModel, na=100, ny=5, nT=25), simulates nsims agents per
model forward, and the per-model summaries and per-period simulation moments are all-gathered once over NCCL.
Weak scaling: `per_gpu` models per rank (8192 = 65536 / 8).  Timed between barriers, max over ranks.
usage: torchrun --nproc-per-node N tools/c5_sweep.py [per_gpu] [nsims]"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dcegm_jl_b200 import Param, _abi, sweep  # noqa: E402
from dcegm_jl_b200.param import params_array  # noqa: E402
from dcegm_jl_b200.solver import SUMMARY_DTYPE, Solver  # noqa: E402

per = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
nsims = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = per * world
rng = np.random.default_rng(25032020 + rank)  # SURVEY.md §8(d) C5 ranges; every rank draws its own shard
base = Param(par=dict(na=100, nsims=nsims)).to_c()
arr = params_array([base] * per)
for i in range(per):
    a = arr[i]
    a.beta, a.R, a.alpha = float(rng.uniform(0.5, 1)), float(rng.uniform(1, 1.5)), float(rng.uniform(0, 0.5))
    a.sigma, a.rho, a.pension = float(rng.uniform(0, 0.35)), float(rng.uniform(0, 0.95)), float(rng.uniform(0, 1))
    a.lambda_ = float(10 ** rng.uniform(np.log10(2e-7), 0))
dev = torch.device("cuda", local)
nT = 25
rd = sweep.record_dtype(nT)
rec = rd.itemsize  # dcegm_summary + simulation status + per-period moments: ONE all-gather per sweep
send = torch.empty(rec * per, dtype=torch.uint8, device=dev)


def job(s):
    s.solve(_abi.MODEL_G, arr, fetch=False)
    t_solve = s.timing().kernel_ms
    s.simulate()
    s.sweep_records_to_device(send.data_ptr(), rec * per, nT)
    out = sweep.gather_records(send, n, rec, dist if world > 1 else None)
    torch.cuda.synchronize()
    return out, t_solve


with Solver(local) as s:
    job(s)  # warm-up (allocations, NCCL communicator)
    best, t_solve = 1e9, 0.0
    for _ in range(2):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out, t_solve = job(s)
        if world > 1:
            dist.barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        best = min(best, float(dt.item()))
    recs = np.frombuffer(out.cpu().numpy().tobytes(), dtype=rd)
    st = recs["summary"]["status"].astype(np.int64)
    ok = int(((st & _abi.ST_FATAL_MASK) == 0).sum())  # only models without a fatal status bit count as solved
    if rank == 0:
        names = {b: k for k, b in vars(_abi).items() if k.startswith("ST_") and isinstance(b, int) and b and (b & (b - 1)) == 0}
        print(json.dumps({"config": f"C5 sweep: {n} GModels (na=100, ny=5, nT=25) + {nsims} agents each on {world} GPU(s)",
                          "n_gpus": world, "models": n, "solved": ok, "seconds": best, "models_per_s": ok / best,
                          "agent_periods_per_s_all_models": n * nsims * nT / best, "solve_kernel_ms_rank0": t_solve,
                          "allgather_bytes": int(out.numel()), "collectives_per_sweep": 1,
                          "status_bits": {names[1 << b]: int(((st >> b) & 1).sum()) for b in range(16) if ((st >> b) & 1).any()},
                          "masked_moment_models": int(np.isnan(recs["moments"][:, 0, 2]).sum())}))
if world > 1:
    dist.destroy_process_group()
