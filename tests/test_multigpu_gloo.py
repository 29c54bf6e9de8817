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
