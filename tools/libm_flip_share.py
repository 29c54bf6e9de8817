"""Share of C3 models whose DISCRETE structure (line lengths, sections, intersections, removed sets, status) differs
between the oracle's two arithmetic modes: libm pow/exp/log (the mode pinned to the MATLAB golden files) and dc_math.h
(the mode the GPU reproduces bit for bit).  SURVEY.md App. B.4 expected none.  CPU only.
usage: libm_flip_share.py [nmodels]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench_configs import c3_params  # noqa: E402
from dcegm_jl_b200 import _abi  # noqa: E402
from dcegm_jl_b200.param import params_array  # noqa: E402
from oracle import oracle  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
arr = c3_params(n)
thr = os.cpu_count() or 1
flips, worst, status_diff, examples = 0, 0.0, 0, []
for c0 in range(0, n, 256):
    sub = params_array([arr[i] for i in range(c0, min(n, c0 + 256))])
    a = oracle.solve(_abi.MODEL_F, sub, nthreads=thr, portable=True)
    b = oracle.solve(_abi.MODEL_F, sub, nthreads=thr, portable=False)
    for m in range(len(sub)):
        same = (np.array_equal(a.lens[m], b.lens[m]) and np.array_equal(a.meta[m], b.meta[m])
                and np.array_equal(a.removed[m], b.removed[m]) and a.status[m] == b.status[m])
        status_diff += int(a.status[m] != b.status[m])
        if not same:
            flips += 1
            if len(examples) < 8:
                it = int(np.argwhere((a.lens[m] != b.lens[m]).any(axis=(1, 2, 3)) | (a.meta[m] != b.meta[m]).any(axis=(1, 2))).max())
                examples.append(dict(model=c0 + m, sigma=sub[m].sigma, lam=sub[m].lambda_, last_period_that_differs=it + 1,
                                     status=(hex(int(a.status[m])), hex(int(b.status[m])))))
        else:
            k = np.arange(a.lay.cap)
            msk = np.broadcast_to((k[None, None, None, None, :] < a.lens[m][..., None])[..., None], a.lines[m].shape)
            x, y = a.lines[m][msk], b.lines[m][msk]
            ok = np.isfinite(x) & np.isfinite(y)
            worst = max(worst, float(np.max(np.abs(x[ok] - y[ok]) / (1e-2 + np.abs(y[ok])))))
print(json.dumps(dict(models=n, structure_differs=flips, share=flips / n, status_differs=status_diff,
                      max_rel_dev_where_structure_equal=worst, examples=examples)))
