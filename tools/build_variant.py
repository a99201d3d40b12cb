#!/usr/bin/env python
"""Build a variant of libpolar_cuda.so with extra nvcc flags (A/B experiments): build_variants/NAME/libpolar_cuda.so, loaded
by POLAR_CUDA_LIB=build_variants/NAME/libpolar_cuda.so.  usage: tools/build_variant.py NAME -DFOO=1 [...]"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from antpolarcodes_b200 import build as b  # noqa: E402

name, extra = sys.argv[1], sys.argv[2:]
out = os.path.join(ROOT, "build_variants", name)
os.makedirs(out, exist_ok=True)


def one(src):
    obj = os.path.join(out, os.path.splitext(src)[0] + ".o")
    subprocess.check_call(["/usr/local/cuda/bin/nvcc"] + b.NVCC_FLAGS + extra + ["-c", os.path.join(b.CSRC, src), "-o", obj], cwd=b.CSRC,
                          stderr=subprocess.DEVNULL)
    return obj


with ThreadPoolExecutor(8) as ex:
    objs = list(ex.map(one, b.SOURCES))
lib = os.path.join(out, "libpolar_cuda.so")
subprocess.check_call(["/usr/local/cuda/bin/nvcc", "--shared", "-cudart", "shared", "-o", lib] + objs, stderr=subprocess.DEVNULL)
for o in objs:
    os.remove(o)
print(lib)
