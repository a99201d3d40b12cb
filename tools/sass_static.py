#!/usr/bin/env python
"""Static SASS instruction count per inlined source function / line of one kernel.
usage: sass_static.py <kernel-mangled-substring> [libpolar_cuda.so]"""
import collections, os, re, subprocess, sys, tempfile
ksub = sys.argv[1]
so = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(__file__), "..", "antpolarcodes_b200", "libpolar_cuda.so")
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
inside, cur = False, None
byline = collections.Counter()
for ln in dis.splitlines():
    if ln.startswith(".text."):
        inside = ksub in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
        byline[cur] += 1
src = {}
def func_of(f, l):
    path = os.path.join(os.path.dirname(__file__), "..", "antpolarcodes_b200", "csrc", f)
    if not os.path.exists(path):
        return f
    if f not in src:
        src[f] = open(path).read().splitlines()
    for i in range(l - 1, -1, -1):
        m = re.match(r"^(?:POLAR_DEV|__global__|template|static|inline)?.*?\b(\w+)\s*\([^;]*$", src[f][i])
        if m and not src[f][i].startswith((" ", "\t", "#", "}", "/")):
            return m.group(1)
    return f
byfunc = collections.Counter()
for (k, n) in byline.items():
    if k:
        byfunc[func_of(*k)] += n
print("total", sum(byline.values()))
for k, n in byfunc.most_common(25):
    print("%6d  %s" % (n, k))
