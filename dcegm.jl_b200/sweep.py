"""Multi-GPU parameter sweeps: contiguous sharding of a batch of independent parameterisations over the
ranks of one torch.distributed job, and the ONE collective of a sweep — an all-gather of fixed-size per-model
records (include/dcegm.h: dcegm_sweep_records_to_device = dcegm_summary + simulation status + per-period
simulation moments).  SURVEY.md §8(e).

The reference has no counterpart (single-threaded Julia); models are independent, so there is no data-path
collective.  A sweep may mix grid sizes (north_star: "parameterisations (lambda, sigma, grid sizes)"): every rank
solves its shard group by group, one launch geometry (na, ny, nT) per group.  On GPUs the gather runs over NCCL on
device buffers filled by the library; the same code runs on the gloo backend with host tensors and a stand-in
backend (tests/test_multigpu_gloo.py)."""
import numpy as np

from . import _abi


def shard_range(nmodels, rank, world):
    """models [lo, hi) of `rank`: contiguous blocks, the first nmodels % world ranks get one more"""
    if not (0 <= rank < world) or nmodels < 0:
        raise ValueError("bad rank/world/nmodels")
    base, extra = divmod(nmodels, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_counts(nmodels, world):
    return [shard_range(nmodels, r, world)[1] - shard_range(nmodels, r, world)[0] for r in range(world)]


def gather_records(local_bytes, nmodels, record_size, dist=None):
    """all-gather the per-rank record buffers (uint8 torch tensors of counts[rank]*record_size bytes) into the
    global array in rank order.  Ragged (and empty) shards are padded to the largest shard so that a single
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


gather_summaries = gather_records  # the name used before records also carried the simulation moments


def summaries_from_bytes(t):
    from .solver import SUMMARY_DTYPE

    return np.frombuffer(t.cpu().numpy().tobytes(), dtype=SUMMARY_DTYPE)


def record_dtype(nT_pad):
    """numpy mirror of the record written by dcegm_sweep_records_to_device"""
    from .solver import SUMMARY_DTYPE

    return np.dtype([("summary", SUMMARY_DTYPE), ("sim_status", np.int32), ("nT", np.int32),
                     ("moments", np.float64, (nT_pad, _abi.SIM_NMOM))])


def geometry_groups(params, lo, hi):
    """indices of params[lo:hi] grouped by launch geometry (na, ny, nT), groups in order of first appearance,
    members in batch order — every rank derives every other rank's grouping from the parameter list alone"""
    groups = {}
    for i in range(lo, hi):
        p = params[i]
        groups.setdefault((int(p.na), int(p.ny), int(p.nT)), []).append(i)
    return list(groups.values())


class GpuBackend:
    """solves / simulates one geometry group on this rank's GPU and returns its records as a device tensor"""

    def __init__(self, device):
        import torch

        from .solver import Solver

        self.torch = torch
        self.dev = torch.device("cuda", device)
        self.solver = Solver(device)

    def close(self):
        self.solver.close()

    def empty(self, nbytes):
        return self.torch.empty(nbytes, dtype=self.torch.uint8, device=self.dev)

    def records(self, model, plist, simulate, keep_lines, nT_pad, out):
        from .param import params_array

        s = self.solver
        s.solve(model, params_array(plist), summary_only=not (keep_lines or simulate), fetch=False)
        if simulate:
            s.simulate()
        s.sweep_records_to_device(out.data_ptr(), out.numel(), nT_pad)


def run_sweep(model, params, simulate=False, dist=None, device=0, keep_lines=False, backend=None):
    """One parameter sweep on the ranks of `dist` (or one GPU): rank r solves the contiguous shard
    shard_range(len(params), r, world) of the batch on its GPU — one launch per (na, ny, nT) group —, optionally
    simulates it forward (simulate.jl:61-203), and the per-model records (summary + per-period simulation moments,
    NaN for status-flagged models) are all-gathered ONCE.  Returns (summaries [n], moments [n, max nT, 8] or None)
    in global model order on every rank.  A rank whose shard is empty takes part in the collective only."""
    import torch

    world = dist.get_world_size() if dist is not None and dist.is_initialized() else 1
    rank = dist.get_rank() if world > 1 else 0
    n = len(params)
    nT_pad = max((int(p.nT) for p in params), default=1)
    rd = record_dtype(nT_pad)
    rec = rd.itemsize
    lo, hi = shard_range(n, rank, world)
    own = backend is None
    if own and hi > lo:
        backend = GpuBackend(device)
    try:
        if hi > lo:
            send = backend.empty(rec * (hi - lo))
            at = 0
            for idx in geometry_groups(params, lo, hi):  # the send buffer holds the shard group by group
                backend.records(model, [params[i] for i in idx], simulate, keep_lines, nT_pad, send[at * rec:(at + len(idx)) * rec])
                at += len(idx)
        else:
            send = torch.zeros(0, dtype=torch.uint8, device=torch.device("cuda", device) if own else "cpu")
        out = gather_records(send, n, rec, dist if world > 1 else None)
    finally:
        if own and backend is not None:
            backend.close()
    recs = np.frombuffer(out.cpu().numpy().tobytes(), dtype=rd)
    # undo the grouping: rank r's part lists its shard group by group
    order = [i for r in range(world) for idx in geometry_groups(params, *shard_range(n, r, world)) for i in idx]
    inv = np.empty(n, np.int64)
    inv[np.asarray(order, np.int64)] = np.arange(n)
    recs = recs[inv]
    return recs["summary"].copy(), (recs["moments"].copy() if simulate else None)
