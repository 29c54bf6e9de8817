"""Packs the reference's own golden data into one fixture file.

Source: /root/reference/matlab/output/{policy,value}_{1,2}_{1..25}.csv — 100 two-row (x;y) CSV
files with 8 significant digits written by matlab/bench.m:13-26 through matlab/polyline.m:472-484
and compared by test/F_test.jl:47-69 under na=500, beta=1/1.05, sigma=0.35, R=1.05, lambda=2e-6,
delta=0 (F_test.jl:34-40).  The data are stored verbatim (float64 of the printed decimals).

Run in the build container (the reference tree does not exist on the GPU box):
    python tests/golden/make_golden.py
"""
import os

import numpy as np

SRC = "/root/reference/matlab/output"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "matlab_retirement_na500.npz")

out = {}
for kind in ("policy", "value"):
    for i in (1, 2):
        for it in range(1, 26):
            a = np.loadtxt(os.path.join(SRC, f"{kind}_{i}_{it}.csv"), delimiter=",")
            out[f"{kind}_{i}_{it}"] = a
np.savez_compressed(DST, **out)
print(DST, os.path.getsize(DST), "bytes,", len(out), "arrays")
