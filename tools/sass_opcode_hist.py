"""Dynamic opcode-class histogram of one kernel from an `ncu --page source --csv` export (SASS view with the
per-instruction "Instructions Executed" and stall samples), and the shared-memory bank-conflict excess.
usage: sass_opcode_hist.py <source.csv> [topN]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 28
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
col = {h: i for i, h in enumerate(rows[hi])}
CLASS = [("FP64 (DFMA/DADD/DMUL/DSETP)", r"^D(FMA|ADD|MUL|SETP)"), ("MUFU", r"^MUFU"), ("shared load (LDS)", r"^LDS"),
         ("shared store (STS)", r"^STS"), ("global/local ld/st", r"^(LDG|STG|LDL|STL|LD$|ST$|ATOM|RED)"),
         ("constant/uniform load", r"^(LDC|LDCU|ULDC)"), ("select (FSEL/SEL)", r"^(FSEL|SEL)"),
         ("int/move (IMAD/IADD3/VIADD/LEA/LOP3/MOV/SHF/VIMNMX/...)", r"^(IMAD|IADD|VIADD|LEA|LOP3|MOV|SHF|VIMNMX|IMNMX|PRMT|I2F|F2I|IABS|FLO|POPC|BREV|CS2R|S2R)"),
         ("compare/predicate (ISETP/FSETP/PLOP3)", r"^(ISETP|FSETP|PLOP3|P2R|R2P)"), ("uniform datapath (U*)", r"^(U[A-Z]|R2UR|S2UR)"),
         ("branch/call/sync (BRA/CALL/RET/BSSY/BSYNC/BAR/WARPSYNC)", r"^(BRA|CALL|RET|BSSY|BSYNC|BAR|WARPSYNC|EXIT|BRX|JMP|NOP|YIELD)"),
         ("shuffle/vote", r"^(SHFL|VOTE|MATCH|REDUX)")]
inst = collections.Counter()
samp = collections.Counter()
ops = collections.Counter()
tot_i = tot_s = 0
wave = ideal = 0
for r in rows[hi + 1:]:
    src = re.sub(r"^@!?U?P\d+\s+", "", r[col["Source"]].strip())
    op = src.split()[0].split(".")[0] if src else "?"
    n = int(float(r[col["Instructions Executed"]] or 0))
    s = int(float(r[col["# Samples"]] or 0))
    k = next((name for name, pat in CLASS if re.match(pat, op)), "other: " + op)
    inst[k] += n
    samp[k] += s
    ops[op] += n
    tot_i += n
    tot_s += s
    try:
        wave += int(float(r[col["L1 Wavefronts Shared"]] or 0))
        ideal += int(float(r[col["L1 Wavefronts Shared Ideal"]] or 0))
    except ValueError:
        pass
print(f"warp instructions executed {tot_i}, stall samples {tot_s}")
for k, n in inst.most_common(topn):
    print(f"{k:72s} inst {100.0 * n / tot_i:6.2f}%  samples {100.0 * samp[k] / max(tot_s, 1):6.2f}%")
print("top opcodes:", ", ".join(f"{o} {100.0 * n / tot_i:.1f}%" for o, n in ops.most_common(16)))
if ideal:
    print(f"shared-memory wavefronts {wave} vs ideal {ideal}: {wave / ideal:.2f}x (bank conflicts)")
