"""Multi-GPU parameter sweeps: contiguous sharding of a batch of independent parameterisations over the
ranks of one torch.distributed job, and the one collective of a sweep — an all-gather of the fixed-size
per-model summary records (include/dcegm.h: dcegm_summary).  SURVEY.md §8(e).

The reference has no counterpart (single-threaded Julia); models are independent, so there is no
data-path collective.  On GPUs the gather runs over NCCL on device buffers filled by
dcegm_summaries_to_device; the same code runs on the gloo backend with host tensors (tests)."""
import numpy as np


def shard_range(nmodels, rank, world):
    """models [lo, hi) of `rank`: contiguous blocks, the first nmodels % world ranks get one more"""
    if not (0 <= rank < world) or nmodels < 0:
        raise ValueError("bad rank/world/nmodels")
    base, extra = divmod(nmodels, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_counts(nmodels, world):
    return [shard_range(nmodels, r, world)[1] - shard_range(nmodels, r, world)[0] for r in range(world)]


def gather_summaries(local_bytes, nmodels, record_size, dist=None, device=None):
    """all-gather the per-rank summary buffers (uint8 torch tensors of counts[rank]*record_size bytes)
    into the global array in model order.  Ragged shards are padded to the largest shard so that a single
    all_gather_into_tensor suffices.  Returns a uint8 tensor of nmodels*record_size bytes."""
    import torch

    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local_bytes[: nmodels * record_size]
    world, rank = dist.get_world_size(), dist.get_rank()
    counts = shard_counts(nmodels, world)
    mx = max(counts) * record_size
    send = local_bytes
    if send.numel() != mx:
        send = torch.zeros(mx, dtype=torch.uint8, device=local_bytes.device)
        send[: counts[rank] * record_size] = local_bytes[: counts[rank] * record_size]
    recv = torch.empty(mx * world, dtype=torch.uint8, device=send.device)
    dist.all_gather_into_tensor(recv, send)
    if all(c == counts[0] for c in counts):
        return recv
    parts = [recv[r * mx: r * mx + counts[r] * record_size] for r in range(world)]
    return torch.cat(parts)


def summaries_from_bytes(t):
    from .solver import SUMMARY_DTYPE

    return np.frombuffer(t.cpu().numpy().tobytes(), dtype=SUMMARY_DTYPE)


def run_sweep(model, params, simulate=False, dist=None, device=0, keep_lines=False):
    """One parameter sweep on the ranks of `dist` (or one GPU): rank r solves the contiguous shard
    shard_range(len(params), r, world) of the batch on its GPU, optionally simulates it forward
    (simulate.jl:61-203), and the per-model summary records — plus the per-period simulation moments — are
    all-gathered once over NCCL from device buffers.  Returns (summaries [n], moments [n,nT,8] or None)
    in global model order on every rank."""
    import torch

    from . import _abi
    from .param import params_array
    from .solver import SUMMARY_DTYPE, Solver

    world = dist.get_world_size() if dist is not None and dist.is_initialized() else 1
    rank = dist.get_rank() if world > 1 else 0
    n = len(params)
    lo, hi = shard_range(n, rank, world)
    arr = params_array(list(params[lo:hi]))
    cnt = hi - lo
    dev = torch.device("cuda", device)
    with Solver(device) as s:
        s.solve(model, arr, summary_only=not (keep_lines or simulate), fetch=False)
        nT = s.layout().nT
        sb = SUMMARY_DTYPE.itemsize * cnt
        send = torch.empty(sb, dtype=torch.uint8, device=dev)
        s.summaries_to_device(send.data_ptr(), sb)
        out = gather_summaries(send, n, SUMMARY_DTYPE.itemsize, dist if world > 1 else None)
        summ = summaries_from_bytes(out)
        mom = None
        if simulate:
            s.simulate()
            rec = nT * _abi.SIM_NMOM * 8
            sendm = torch.empty(rec * cnt, dtype=torch.uint8, device=dev)
            s.sim_moments_to_device(sendm.data_ptr(), rec * cnt)
            outm = gather_summaries(sendm, n, rec, dist if world > 1 else None)
            mom = np.frombuffer(outm.cpu().numpy().tobytes(), dtype=np.float64).reshape(n, nT, _abi.SIM_NMOM)
    return summ, mom
