"""Aggregate an ncu SASS source page (csv) by the OUTERMOST inlining site inside one source file
(nvdisasm -g -c listing), i.e. by the line of the kernel body that the code was inlined into.
usage: ncu_by_phase.py <ncu --page source --csv> <nvdisasm -g -c listing> <kernel name> <file substring> [line ranges a-b:name ...]"""
import collections
import csv
import re
import sys

src_csv, dis, kname, fsub = sys.argv[1:5]
ranges = []
for a in sys.argv[5:]:
    r, name = a.split(":")
    lo, hi = r.split("-")
    ranges.append((int(lo), int(hi), name))
rows = list(csv.reader(open(src_csv)))
hi_ = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
col = {h: i for i, h in enumerate(rows[hi_])}
sass = rows[hi_ + 1:]
lines = open(dis).read().split("\n")
start = next(i for i, l in enumerate(lines) if ".text." in l and kname in l)
tags = []
chain = []
last_outer = None
for l in lines[start + 1:]:
    if ".text." in l and kname not in l:
        break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
    if m:
        chain = [(m.group(1), int(m.group(2)))]
        if m.group(3):
            chain.append((m.group(3), int(m.group(4))))
        continue
    m2 = re.match(r'\s*//## +inlined at "([^"]+)", line (\d+)', l)
    if m2:
        chain.append((m2.group(1), int(m2.group(2))))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+\S", l):
        # nvdisasm prints only the innermost location: carry the last line seen in the kernel's own
        # file forward (code is laid out roughly in program order), a heuristic attribution
        for f, ln in chain:
            if fsub in f:
                last_outer = ln
        tags.append(last_outer)
print("sass rows", len(sass), "tagged", len(tags))
REASONS = ["stall_wait", "stall_no_inst", "stall_barrier", "stall_short_sb", "stall_long_sb", "stall_math",
           "stall_branch_resolving", "stall_not_selected", "stall_selected"]


def num(r, name):
    v = r[col[name]]
    return int(float(v)) if v else 0


agg = collections.defaultdict(lambda: [0] * (3 + len(REASONS)))
tot = [0] * (3 + len(REASONS))
for r, t in zip(sass, tags):
    key = t
    if ranges:
        key = "other"
        for lo, hi, name in ranges:
            if t is not None and lo <= t <= hi:
                key = name
    vals = [num(r, "Instructions Executed"), num(r, "# Samples"), 1] + [num(r, k) for k in REASONS]
    for i, v in enumerate(vals):
        agg[key][i] += v
        tot[i] += v
print("total inst", tot[0], "samples", tot[1])
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:60]:
    s = f"{str(k):>14}  inst {100 * v[0] / tot[0]:5.2f}%  samples {100 * v[1] / tot[1]:5.2f}%  sass {v[2]:5d} |"
    for i, kk in enumerate(REASONS[:7]):
        s += f" {kk[6:]} {100 * v[3 + i] / max(v[1], 1):4.0f}%"
    print(s)
