"""Join an ncu SASS-level source page (csv) with nvdisasm -g line info and aggregate executed
instructions / stall samples per CUDA source line (with the main stall reasons).
usage: ncu_by_line.py <ncu --page source --csv> <nvdisasm -g -c listing> <kernel name> [topN]"""
import collections
import csv
import re
import sys

src_csv, dis, kname = sys.argv[1], sys.argv[2], sys.argv[3]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
rows = list(csv.reader(open(src_csv)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
col = {h: i for i, h in enumerate(rows[hi])}
sass = rows[hi + 1:]
lines = open(dis).read().split("\n")
start = next(i for i, l in enumerate(lines) if ".text." in l and kname in l)
cur = ("?", 0)
tags = []
for l in lines[start:]:
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        # inlined code: attribute to the innermost line, remember the outermost call site
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+\S", l):
        tags.append(cur)
    if len(tags) >= len(sass):
        break
print("sass rows", len(sass), "tagged", len(tags))
REASONS = ["stall_wait", "stall_no_inst", "stall_barrier", "stall_short_sb", "stall_long_sb", "stall_math",
           "stall_branch_resolving", "stall_not_selected", "stall_selected"]
NCOL = 3 + len(REASONS) + 2  # ... + thread instructions, predicated-on thread instructions
agg = collections.defaultdict(lambda: [0] * NCOL)
tot = [0] * NCOL


def num(r, name):
    v = r[col[name]]
    return int(float(v)) if v else 0


for r, t in zip(sass, tags):
    vals = ([num(r, "Instructions Executed"), num(r, "# Samples"), 1] + [num(r, k) for k in REASONS]
            + [num(r, "Thread Instructions Executed"), num(r, "Predicated-On Thread Instructions Executed")])
    for i, v in enumerate(vals):
        agg[t][i] += v
        tot[i] += v
print("total inst", tot[0], "samples", tot[1])
print("stall mix: " + "  ".join(f"{k[6:]} {100 * tot[3 + i] / tot[1]:.1f}%" for i, k in enumerate(REASONS)))
print(f"warp divergence: {tot[-2] / max(tot[0], 1):.2f} active threads per warp instruction, "
      f"{tot[-1] / max(tot[0], 1):.2f} predicated-on")
byfile = collections.defaultdict(lambda: [0] * NCOL)
for (f, ln), v in agg.items():
    for i in range(len(v)):
        byfile[f][i] += v[i]


def fmt(v):
    s = f"inst {100 * v[0] / tot[0]:5.2f}%  samples {100 * v[1] / tot[1]:5.2f}%  sass {v[2]:5d} |"
    for i, k in enumerate(REASONS[:6]):
        s += f" {k[6:]} {100 * v[3 + i] / max(v[1], 1):4.0f}%"
    s += f" | thr/inst {v[-2] / max(v[0], 1):4.1f} on {v[-1] / max(v[0], 1):4.1f}"
    return s


for f, v in sorted(byfile.items(), key=lambda kv: -kv[1][1]):
    print(f"{f:28s} {fmt(v)}")
print("--- top lines by samples")
for (f, ln), v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:topn]:
    print(f"{f + ':' + str(ln):28s} {fmt(v)}")
