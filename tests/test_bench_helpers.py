"""bench.py pieces that need no GPU: the synthetic draws are reproducible, and the ncu counters the bench line quotes
(profiles/solve_f_counters.json) were captured on exactly the kernel sources in the tree — bench.py prints null for
`roofline.traffic` / `roofline_fp64.achieved` otherwise, so a kernel change without a re-capture fails here first."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_counters_file_matches_the_kernel_sources():
    import bench

    c, why = bench.ncu_counters()
    assert c is not None, why
    assert c["source_sha16"] == bench.kernel_source_hash()
    assert c["models"] > 0 and 0.9 < (c["dram_bytes_read"] + c["dram_bytes_write"]) / c["models"] / bench.ALGO_BYTES_PER_MODEL < 1.1
    assert 0.2 < c["fp64_warp_inst"] / c["warp_inst"] < 0.5


def test_draws_are_reproducible_and_shards_are_slices_of_one_draw():
    from bench_configs import c2_params, c3_params, c5_params

    a = c3_params(8)
    b = c3_params(4, offset=4)
    assert [x.lambda_ for x in a][4:] == [x.lambda_ for x in b] and [x.sigma for x in a][4:] == [x.sigma for x in b]
    assert all(2e-6 <= x.lambda_ <= 0.2 and 0 <= x.sigma <= 0.6 and x.na == 500 and x.nT == 25 for x in a)
    assert [x.beta for x in c2_params(5)] == [x.beta for x in c2_params(5)]
    r0, r1 = c5_params(3, 10, rank=0), c5_params(3, 10, rank=1)
    assert [x.rho for x in r0] != [x.rho for x in r1] and all(0 <= x.rho < 0.95 and x.nsims == 10 for x in r0)


def test_status_counts_ignores_informational_bits():
    from bench_configs import status_counts
    from dcegm_jl_b200 import _abi

    ok, bits = status_counts(np.array([0, _abi.ST_EXTRA_RIGHT, _abi.ST_NONMONO_FIRST, _abi.ST_DUP_KNOT, _abi.ST_DUP_KNOT | _abi.ST_EXTRA_RIGHT], np.int32))
    assert ok == 3 and bits == {"ST_EXTRA_RIGHT": 2, "ST_NONMONO_FIRST": 1, "ST_DUP_KNOT": 2}
