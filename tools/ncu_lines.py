"""Per-CUDA-source-line totals from `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`.
usage: ncu_lines.py <csv> [topN] [sort: inst|samples]"""
import collections
import csv
import sys

path = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
key = sys.argv[3] if len(sys.argv) > 3 else "inst"
cur_file, col = None, None
agg = collections.defaultdict(lambda: [0, 0, 0, ""])
REASONS = ["stall_wait", "stall_no_inst", "stall_barrier", "stall_short_sb", "stall_long_sb", "stall_math", "stall_branch_resolving"]
stall = collections.defaultdict(lambda: [0] * len(REASONS))
for r in csv.reader(open(path)):
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        col = {}
        for i, h in enumerate(r):
            col.setdefault(h, i)
        continue
    if r[0] == "Function Name" or col is None:
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    if not r[col["Address"]]:
        if r[1].strip():
            agg[(cur_file, ln)][3] = r[1].strip()
        continue

    def num(name):
        v = r[col[name]]
        return int(float(v)) if v and v != "-" else 0

    a = agg[(cur_file, ln)]
    a[0] += num("Instructions Executed")
    a[1] += num("# Samples")
    a[2] += 1
    for i, k in enumerate(REASONS):
        stall[(cur_file, ln)][i] += num(k)
ti = sum(v[0] for v in agg.values())
ts = sum(v[1] for v in agg.values())
print(f"total inst {ti} samples {ts}")
byfile = collections.defaultdict(lambda: [0, 0, 0])
for (f, ln), v in agg.items():
    for i in range(3):
        byfile[f][i] += v[i]
for f, v in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print(f"{f:28s} inst {100 * v[0] / ti:5.2f}%  samples {100 * v[1] / ts:5.2f}%  sass {v[2]}")
idx = 0 if key == "inst" else 1
for (f, ln), v in sorted(agg.items(), key=lambda kv: -kv[1][idx])[:topn]:
    st = stall[(f, ln)]
    tops = sorted(zip(st, REASONS), reverse=True)[:2]
    ss = " ".join(f"{k[6:]} {100 * s / max(v[1], 1):.0f}%" for s, k in tops)
    print(f"{f}:{ln:<5d} inst {100 * v[0] / ti:5.2f}% smp {100 * v[1] / ts:5.2f}% sass {v[2]:4d} [{ss}] | {v[3][:90]}")
