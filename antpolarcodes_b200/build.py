"""Builds antpolarcodes_b200/libpolar_cuda.so (sm_100a) in-tree with nvcc.  Used by __graft_entry__.build().
One translation unit per kernel family, compiled in parallel into build/*.o and linked into the shared library."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
OUT = os.path.join(HERE, "libpolar_cuda.so")
SOURCES = ["polar_cuda.cu", "scl_i8.cu", "scl_f32.cu", "fastssc_i8.cu", "fastssc_f32.cu", "schedule.cpp", "crc_host.cpp", "construct.cpp"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "-Xcompiler", "-Wno-unused-function"]


def _deps():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC) if not os.path.isdir(os.path.join(CSRC, f))] + \
           [os.path.join(HERE, "..", "include", "polar_cuda.h")]


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(d) > t for d in _deps())


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(OBJ, exist_ok=True)
    newest_hdr = max(os.path.getmtime(d) for d in _deps() if not d.endswith((".cu", ".cpp")))

    def compile_one(src):
        obj = os.path.join(OBJ, os.path.splitext(src)[0] + ".o")
        path = os.path.join(CSRC, src)
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(path), newest_hdr):
            return obj
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
        subprocess.check_call(cmd, cwd=CSRC)
        return obj

    with ThreadPoolExecutor(max(1, min(len(SOURCES), os.cpu_count() or 1))) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    subprocess.check_call([nvcc, "--shared", "-cudart", "shared", "-o", OUT + ".tmp"] + objs, cwd=CSRC)
    os.replace(OUT + ".tmp", OUT)  # never leave a half-written library where a snapshot or a loader can see it
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(OUT)
