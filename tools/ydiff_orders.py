"""Offline study for the compact line format: bytes of the ordinate payload when blocks of 32 points carry
k-th differences of the 64-bit patterns, k = 1..5, and when every block picks its own best k.
Solves a sample of C3 draws with the CPU oracle (portable arithmetic = the kernels' bits).
    python tools/ydiff_orders.py [nmodels]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle  # noqa: E402
from bench_configs import c3_params  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
arr = c3_params(4096)
step = 4096 // n
ps = [arr[i * step] for i in range(n)]
sol = oracle.solve_f(ps, portable=True, nthreads=8)
lens = sol.lens.reshape(-1)
lines = sol.lines.reshape(len(lens), -1, 2)


def widths(R):  # bytes of the shortest two's complement form of every residual
    r = R.view(np.int64)
    mag = np.where(r < 0, ~r, r).astype(np.uint64)
    bits = np.where(mag == 0, 0, np.floor(np.log2(np.maximum(mag, 1).astype(np.float64))).astype(np.int64) + 1)
    # log2 on float64 is inexact near 2^53+: fix by checking
    bits = np.where((mag >> np.minimum(bits, 63).astype(np.uint64)) > 0, bits + 1, bits)
    bits = np.where((bits > 0) & ((mag >> np.maximum(bits - 1, 0).astype(np.uint64)) == 0), bits - 1, bits)
    return np.where(r == 0, 0, (bits + 1 + 7) // 8)


tot = np.zeros(6, np.int64)      # payload bytes per fixed order 1..5 (index 0 unused)
best = 0
choice = np.zeros(6, np.int64)
npts = 0
with np.errstate(over="ignore"):
    for s in range(len(lens)):
        m = int(lens[s])
        if m < 3:
            continue
        I = np.ascontiguousarray(lines[s, :m, 1]).view(np.uint64)
        npts += m
        D = {0: I}
        for k in range(1, 6):
            d = np.zeros(m, np.uint64)
            d[k:] = D[k - 1][k:] - D[k - 1][k - 1:-1] if k > 1 else I[1:] - I[:-1]
            D[k] = d
        W = {k: widths(D[k]) for k in range(1, 6)}
        for b0 in range(0, m, 32):
            cnt = min(32, m - b0)
            per = []
            for k in range(1, 6):
                lo = max(b0, 2)  # the first two ordinates travel raw, as today
                w = int(W[k][lo:b0 + cnt].max()) if lo < b0 + cnt else 0
                # a k-th difference at position j < k is not defined from in-line data: those positions fall back
                # to a lower order in the real format; ignored here (first block only)
                tot[k] += w * cnt
                per.append(w * cnt)
            best += min(per)
            choice[1 + int(np.argmin(per))] += 1
print(f"{n} models, {npts} points in lines of >= 3 points")
for k in range(1, 6):
    print(f"order {k}: {tot[k] / npts:.3f} payload bytes / point")
print(f"best order per block: {best / npts:.3f} bytes / point; blocks choosing order 1..5: {choice[1:].tolist()}")

# per-point widths (no block maximum): how much of the cost is the block's widest residual
hist = {k: np.zeros(10, np.int64) for k in range(1, 6)}
kinds = {0: np.zeros(2), 1: np.zeros(2)}  # policy / value: (points, order-2 block bytes)
with np.errstate(over="ignore"):
    for s in range(len(lens)):
        m = int(lens[s])
        if m < 3:
            continue
        I = np.ascontiguousarray(lines[s, :m, 1]).view(np.uint64)
        prev = I
        for k in range(1, 6):
            d = np.zeros(m, np.uint64)
            d[k:] = prev[k:] - prev[k - 1:-1]
            prev = d
            w = widths(d)[max(k, 2):]
            hist[k] += np.bincount(w, minlength=10)[:10]
            if k == 2:
                kd = s & 1
                kinds[kd][0] += m
                ww = widths(d)
                for b0 in range(0, m, 32):
                    lo = max(b0, 2)
                    if lo < min(b0 + 32, m):
                        kinds[kd][1] += int(ww[lo:min(b0 + 32, m)].max()) * (min(b0 + 32, m) - b0)
for k in range(1, 6):
    h = hist[k]
    print(f"order {k}: per-point width histogram (bytes 0..8) {h[:9].tolist()}, mean {np.dot(h[:9], np.arange(9)) / h.sum():.3f}")
for kd, nm in ((0, "policy"), (1, "value")):
    print(f"{nm} lines: {kinds[kd][1] / kinds[kd][0]:.3f} payload bytes / point at order 2")

# block size x width granularity (order 2): payload + width-field bytes per point
def bitlen(R):
    r = R.view(np.int64)
    mag = np.where(r < 0, ~r, r).astype(np.uint64)
    out = np.zeros(len(r), np.int64)
    for sh in (32, 16, 8, 4, 2, 1):
        big = (mag >> np.uint64(sh)) > 0
        out += np.where(big, sh, 0)
        mag = np.where(big, mag >> np.uint64(sh), mag)
    out += (mag > 0)
    return np.where(r == 0, 0, out + 1)  # + sign bit

cfg = [(B, g) for B in (8, 16, 32) for g in (1, 4, 8)]
cost = {c: 0.0 for c in cfg}
cost_ad = {c: 0.0 for c in cfg}
with np.errstate(over="ignore"):
    for s in range(len(lens)):
        m = int(lens[s])
        if m < 3:
            continue
        I = np.ascontiguousarray(lines[s, :m, 1]).view(np.uint64)
        d1 = np.zeros(m, np.uint64); d1[1:] = I[1:] - I[:-1]
        d2 = np.zeros(m, np.uint64); d2[2:] = d1[2:] - d1[1:-1]
        d3 = np.zeros(m, np.uint64); d3[3:] = d2[3:] - d2[2:-1]
        b2, b3 = bitlen(d2), bitlen(d3)
        b2[:2] = 0; b3[:3] = b2[:3]
        for (B, g) in cfg:
            nb = (m + B - 1) // B
            pad = nb * B - m
            x2 = np.concatenate([b2, np.zeros(pad, np.int64)]).reshape(nb, B).max(axis=1)
            x3 = np.concatenate([b3, np.zeros(pad, np.int64)]).reshape(nb, B).max(axis=1)
            cnt = np.full(nb, B); cnt[-1] = B - pad
            w2 = (x2 + g - 1) // g * g
            w3 = (x3 + g - 1) // g * g
            field = {1: 7, 4: 5, 8: 4}[g]  # bits of the per-block width field (incl. one order bit when adaptive)
            cost[(B, g)] += (w2 * cnt).sum() / 8 + nb * (field - 1) / 8
            cost_ad[(B, g)] += (np.minimum(w2, w3) * cnt).sum() / 8 + nb * field / 8
for c in cfg:
    print(f"block {c[0]:2d}, width granularity {c[1]} bit: order 2 {cost[c] / npts:.3f} B/pt, best of orders 2/3 {cost_ad[c] / npts:.3f} B/pt")
