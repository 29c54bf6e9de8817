"""Multi-GPU sweep check, run under torchrun (one rank per GPU): every rank solves + simulates its shard
of a C5-style GModel batch; the all-gathered summaries and simulation moments must equal, bit for bit,
what rank 0 computes alone for the whole batch.
usage: torchrun --nproc-per-node N tools/sweep_check.py [nmodels] [nsims]"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dcegm_jl_b200 import Param, _abi, sweep  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
nsims = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rng = np.random.default_rng(25032020)
ps = []
for k in range(n):  # SURVEY.md §8(d) C5 ranges
    d = dict(na=100, nsims=nsims, beta=float(rng.uniform(0.5, 1)), R=float(rng.uniform(1, 1.5)), alpha=float(rng.uniform(0, 0.5)),
             sigma=float(rng.uniform(0, 0.35)), rho=float(rng.uniform(0, 0.95)), pension=float(rng.uniform(0, 1)))
    d["lambda"] = float(10 ** rng.uniform(np.log10(2e-7), 0))
    ps.append(Param(par=d))
dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
summ, mom = sweep.run_sweep(_abi.MODEL_G, ps, simulate=True, dist=dist, device=local)
torch.cuda.synchronize()
dist.barrier()
dt = time.perf_counter() - t0
if dist.get_rank() == 0:
    s1, m1 = sweep.run_sweep(_abi.MODEL_G, ps, simulate=True, dist=None, device=local)
    ok = summ.tobytes() == s1.tobytes() and np.array_equal(np.nan_to_num(mom, nan=-1.0), np.nan_to_num(m1, nan=-1.0))
    print(f"sweep_check world={dist.get_world_size()} models={n} nsims={nsims}: gathered == single-GPU: {ok}; "
          f"{dt * 1e3:.1f} ms incl. host setup; flagged={int((summ['status'] != 0).sum())}")
    assert ok
dist.destroy_process_group()
