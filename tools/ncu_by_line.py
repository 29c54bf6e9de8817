"""Join an ncu SASS-level source page (csv) with nvdisasm -g line info and aggregate executed
instructions / stall samples per CUDA source line.
usage: ncu_by_line.py <ncu --page source --csv> <nvdisasm -g listing> [topN]"""
import csv, re, sys, collections
src_csv, dis, topn = sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40
rows = list(csv.reader(open(src_csv)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
sass = rows[hi + 1:]
# nvdisasm: find the kernel's text section and walk it
lines = open(dis).read().split("\n")
start = next(i for i, l in enumerate(lines) if re.match(r"\s*\.text\._ZN5dcegm14solve_f_kernel", l) or "solve_f_kernel" in l and ".text." in l)
cur = ("?", 0); tags = []
for l in lines[start:]:
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)), m.group(3))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+\S", l):
        tags.append(cur)
    if len(tags) >= len(sass): break
print("sass rows", len(sass), "tagged", len(tags))
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0]
for r, t in zip(sass, tags):
    ie = int(float(r[col["Instructions Executed"]] or 0)); ss = int(float(r[col["# Samples"]] or 0))
    key = (t[0], t[1])
    agg[key][0] += ie; agg[key][1] += ss; agg[key][2] += 1
    tot[0] += ie; tot[1] += ss
print("total inst", tot[0], "samples", tot[1])
byfile = collections.defaultdict(lambda: [0, 0, 0])
for (f, ln), v in agg.items():
    for i in range(3): byfile[f][i] += v[i]
for f, v in sorted(byfile.items(), key=lambda kv: -kv[1][1]):
    print(f"{f:24s} inst {100*v[0]/tot[0]:5.1f}%  samples {100*v[1]/tot[1]:5.1f}%  sass {v[2]}")
print("--- top lines by samples")
for (f, ln), v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:topn]:
    print(f"{f}:{ln:<5d} inst {100*v[0]/tot[0]:5.2f}%  samples {100*v[1]/tot[1]:5.2f}%  sass {v[2]}")
