"""Turns an `ncu --metrics ... --csv` log of ONE solve_f_kernel launch into profiles/solve_f_counters.json, the file
bench.py reads its `roofline.traffic` and `roofline_fp64.achieved` from.  The file records a hash of the kernel's
sources (bench.kernel_source_hash): bench.py refuses the counters when the sources have changed since the capture.
usage: ncu_counters_json.py <metrics.csv> <capture name> [out.json]"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import COUNTERS_FILE, kernel_source_hash  # noqa: E402

path, capture = sys.argv[1], sys.argv[2]
out = sys.argv[3] if len(sys.argv) > 3 else COUNTERS_FILE
rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
col = {h: i for i, h in enumerate(rows[0])}
vals, grid = {}, None
for r in rows[1:]:
    if "solve_f_kernel" not in r[col["Kernel Name"]]:
        continue
    vals[r[col["Metric Name"]]] = float(r[col["Metric Value"]].replace(",", ""))
    grid = int(re.match(r"\((\d+)", r[col["Grid Size"]]).group(1))
unit_ns = vals.get("gpu__time_duration.sum")
rec = {"capture": capture, "kernel": "solve_f_kernel", "models": grid, "source_sha16": kernel_source_hash(),
       "dram_bytes_read": vals["dram__bytes_read.sum"], "dram_bytes_write": vals["dram__bytes_write.sum"],
       "fp64_warp_inst": vals["sm__inst_executed_pipe_fp64.sum"], "warp_inst": vals["smsp__inst_executed.sum"],
       "duration_ns_under_ncu": unit_ns,
       "command": "ncu --metrics sm__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum,gpu__time_duration.sum,"
                  "dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:solve_f_kernel "
                  f"python bench.py --batch {grid} --no-cpu --no-configs"}
json.dump(rec, open(out, "w"), indent=1)
print(json.dumps(rec))
