#!/usr/bin/env python
"""Join an ncu SASS-level source page (ncu -i X.ncu-rep --page source --csv) with nvdisasm -g line info
and print executed warp-instructions per source line / per function of the .cuh files.

usage: ncu_lines.py <rep.ncu-rep> <kernel-mangled-substring> [libpolar_cuda.so]
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

rep, ksub = sys.argv[1], sys.argv[2]
so = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(__file__), "..", "antpolarcodes_b200", "libpolar_cuda.so")
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
dis = ""
for cubin in sorted(f for f in os.listdir(tmp) if f.endswith(".cubin")):  # one cubin per translation unit
    d = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    if ksub in d:
        dis += d
addr2line, cur, inside = {}, None, False
for ln in dis.splitlines():
    if ln.startswith(".text."):
        inside = ksub in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m:
        addr2line[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ia, ie, isrc = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("Source")
istall = hdr.index("# Samples")
data = [r for r in rows[2:] if len(r) > ie and r[ie].isdigit()]
base = int(data[0][ia], 16) if data[0][ia].startswith("0x") else int(data[0][ia])
per_line, samples = collections.Counter(), collections.Counter()
tot = 0
for r in data:
    a = (int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])) - base
    n = int(r[ie])
    tot += n
    per_line[addr2line.get(a)] += n
    samples[addr2line.get(a)] += int(r[istall] or 0)
# function map from the source files: line -> enclosing function name
func_of = {}
for f in set(k[0] for k in per_line if k):
    p = os.path.join(os.path.dirname(__file__), "..", "antpolarcodes_b200", "csrc", f)
    if not os.path.exists(p):
        continue
    name = "?"
    for i, l in enumerate(open(p), 1):
        m = re.match(r"^(?:template.*>\s*)?(?:POLAR_DEV|__global__|inline|static)?\s*[\w:<>\*& ]*?\b(\w+)\(.*", l)
        if m and not l.startswith(" ") and not l.startswith("//") and "(" in l and not l.strip().endswith(";"):
            name = m.group(1)
        func_of[(f, i)] = name
per_func, fsamp = collections.Counter(), collections.Counter()
for k, v in per_line.items():
    per_func[(k[0], func_of.get(k, "?")) if k else None] += v
    fsamp[(k[0], func_of.get(k, "?")) if k else None] += samples[k]
stot = sum(samples.values()) or 1
print("total warp-instructions: %d" % tot)
print("--- per function (instr %, stall-sample %)")
for k, v in per_func.most_common(60):
    print("%6.2f%%  %6.2f%%  %s" % (100.0 * v / tot, 100.0 * fsamp[k] / stot, k))
print("--- top lines")
for k, v in per_line.most_common(120):
    print("%6.2f%%  %6.2f%%  %s" % (100.0 * v / tot, 100.0 * samples[k] / stot, k))
