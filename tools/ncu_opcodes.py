#!/usr/bin/env python
"""Executed warp-instructions per SASS opcode from an ncu report (ncu -i X --page source --csv).
usage: ncu_opcodes.py <rep.ncu-rep> [top]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ie, isrc = hdr.index("Instructions Executed"), hdr.index("Source")
cnt, tot = collections.Counter(), 0
FMA = ("FFMA", "FMUL", "FADD", "IMAD", "HFMA2", "HADD2", "HMUL2")
pipe = collections.Counter()
for r in rows[2:]:
    if len(r) <= ie or not r[ie].isdigit():
        continue
    n = int(r[ie])
    s = r[isrc].strip()
    if s.startswith("@"):
        s = s.split(None, 1)[1] if " " in s else s
    op = s.split()[0] if s else "?"
    base = op.split(".")[0]
    cnt[op if base in ("IMAD", "LOP3", "PRMT") and len(op) < 14 else base] += n
    tot += n
    if base in FMA:
        pipe["fma"] += n
    elif base in ("LDS", "STS", "LDG", "STG", "LD", "ST", "LDC", "LDL", "STL", "ATOMS", "RED"):
        pipe["lsu"] += n
    elif base in ("SHFL", "BAR", "BRA", "BSSY", "BSYNC", "WARPSYNC", "EXIT", "CALL", "RET", "JMP", "BRX", "VOTE", "REDUX", "NOP", "S2R", "CS2R", "MATCH", "BMOV", "NANOSLEEP", "YIELD", "DEPBAR", "ERRBAR", "MEMBAR"):
        pipe["ctl/other"] += n
    elif base.startswith("U") or base.startswith("R2U"):
        pipe["uniform"] += n
    else:
        pipe["alu"] += n
print("total %d" % tot)
for k, v in pipe.most_common():
    print("  pipe %-10s %6.2f%%" % (k, 100.0 * v / tot))
for k, v in cnt.most_common(top):
    print("%-22s %6.2f%%" % (k, 100.0 * v / tot))
