"""World-size-2 (and 3, ragged) gloo runs of the sweep host logic: contiguous sharding and the single
all-gather of per-model summary records (dcegm.jl_b200/sweep.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dcegm_jl_b200 import sweep
from dcegm_jl_b200.solver import SUMMARY_DTYPE


def test_shard_range_partitions():
    for n in (0, 1, 7, 4096, 65536, 65537):
        for w in (1, 2, 3, 8):
            r = [sweep.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
            c = sweep.shard_counts(n, w)
            assert max(c) - min(c) <= 1 and sum(c) == n
    with pytest.raises(ValueError):
        sweep.shard_range(4, 2, 2)


def _fake_summaries(lo, hi):
    s = np.zeros(hi - lo, SUMMARY_DTYPE)
    idx = np.arange(lo, hi)
    s["status"] = idx % 3
    s["tot_points"] = idx * 1000
    s["checksum_c"] = idx * 0.5
    s["v1_probe"][:, 0] = -idx
    return s


def _worker(rank, world, port, nmodels, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sweep.shard_range(nmodels, rank, world)
    local = torch.from_numpy(np.frombuffer(_fake_summaries(lo, hi).tobytes(), np.uint8).copy())
    out = sweep.gather_summaries(local, nmodels, SUMMARY_DTYPE.itemsize, dist)
    got = sweep.summaries_from_bytes(out)
    ok = got.tobytes() == _fake_summaries(0, nmodels).tobytes()
    # max-over-ranks timing reduction used by bench.py
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    q.put((rank, ok, float(t.item())))
    dist.destroy_process_group()


class _FakeBackend:
    """stands in for the GPU library in run_sweep: records computed from the parameters alone, on host tensors;
    like the library it refuses a group of mixed geometry"""

    def __init__(self):
        self.calls = []

    def empty(self, nbytes):
        return torch.zeros(nbytes, dtype=torch.uint8)

    def records(self, model, plist, simulate, keep_lines, nT_pad, out):
        geo = {(p.na, p.ny, p.nT) for p in plist}
        assert len(geo) == 1, "one launch geometry per call"
        self.calls.append((geo.pop(), len(plist)))
        rd = sweep.record_dtype(nT_pad)
        r = np.zeros(len(plist), rd)
        for k, p in enumerate(plist):
            r["summary"]["status"][k] = 0
            r["summary"]["max_len"][k] = p.na
            r["summary"]["checksum_c"][k] = p.sigma
            r["sim_status"][k] = 0 if simulate else -1
            r["nT"][k] = p.nT
            r["moments"][k] = np.nan
            if simulate:
                r["moments"][k, : p.nT, :] = p.sigma * np.arange(p.nT)[:, None]
        out.copy_(torch.from_numpy(np.frombuffer(r.tobytes(), np.uint8).copy()))


def _mixed_params(n):
    from dcegm_jl_b200 import Param

    geos = [(100, 5, 25), (60, 3, 10), (200, 7, 25)]
    out = []
    for i in range(n):
        na, ny, nT = geos[(i * 7 + i // 3) % 3]
        out.append(Param(par=dict(na=na, ny=ny, nT=nT, sigma=0.01 * (i + 1))))
    return out


def _sweep_worker(rank, world, port, nmodels, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ncoll = [0]
    orig = dist.all_gather_into_tensor

    def counting(*a, **k):
        ncoll[0] += 1
        return orig(*a, **k)

    dist.all_gather_into_tensor = counting
    params = _mixed_params(nmodels)
    be = _FakeBackend()
    summ, mom = sweep.run_sweep(1, params, simulate=True, dist=dist, backend=be)
    ok = len(summ) == nmodels and mom.shape == (nmodels, 25, 8)
    for i, p in enumerate(params):  # global model order restored although every rank sent its shard group by group
        ok &= int(summ["max_len"][i]) == p.na and float(summ["checksum_c"][i]) == p.sigma
        ok &= bool(np.all(mom[i, : p.nT, 0] == p.sigma * np.arange(p.nT))) and bool(np.isnan(mom[i, p.nT:]).all())
    lo, hi = sweep.shard_range(nmodels, rank, world)
    ok &= sum(c for _, c in be.calls) == hi - lo
    q.put((rank, bool(ok), float(ncoll[0])))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,nmodels", [(2, 23), (3, 2)])
def test_run_sweep_mixed_geometries_one_collective(world, nmodels):
    """a sweep over grid sizes runs one launch per (na, ny, nT) group and exactly one all-gather; with fewer models
    than ranks the rank with the empty shard only takes part in the collective"""
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_sweep_worker, args=(r, world, port, nmodels, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = [q.get(timeout=180) for _ in ps]
    for p in ps:
        p.join(timeout=60)
    assert sorted(r[0] for r in res) == list(range(world))
    assert all(r[1] for r in res)
    assert all(r[2] == 1.0 for r in res)


def test_run_sweep_single_process_groups():
    params = _mixed_params(11)
    be = _FakeBackend()
    summ, mom = sweep.run_sweep(1, params, simulate=False, backend=be)
    assert mom is None and [int(x) for x in summ["max_len"]] == [p.na for p in params]
    assert len(be.calls) == 3 and sum(c for _, c in be.calls) == 11


@pytest.mark.parametrize("world,nmodels", [(2, 64), (3, 10)])
def test_allgather_of_summaries(world, nmodels):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, nmodels, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in ps]
    for p in ps:
        p.join(timeout=60)
    assert sorted(r[0] for r in res) == list(range(world))
    assert all(r[1] for r in res)
    assert all(r[2] == float(world) for r in res)
