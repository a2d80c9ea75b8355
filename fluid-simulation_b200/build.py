"""Build libflipb200.so (hand-written sm_100a CUDA behind the C ABI of include/flipb200.h).

In-tree build with nvcc; the .so is git-ignored but travels to the GPU box with the snapshot.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
# experiments: FLIPB200_BUILD_SUFFIX=_b FLIPB200_BUILD_DEFS="-DX=1 -DY=2" python build.py  -> libflipb200_b.so beside the product
SUFFIX = os.environ.get("FLIPB200_BUILD_SUFFIX", "")
OBJ = os.path.join(HERE, "build" + SUFFIX)
LIB = os.path.join(HERE, f"libflipb200{SUFFIX}.so")
SOURCES = ["ctx.cu", "particles.cu", "rows.cu", "grid.cu", "pcg.cu", "mg.cu", "mg_h16.cu", "pcg_p32.cu", "step.cu", "comm.cu", "noalloc.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-I", os.path.join(ROOT, "include"), "-I", CSRC,
    "-ccbin", "/usr/bin/g++",
] + os.environ.get("FLIPB200_BUILD_DEFS", "").split()


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "flipb200.h"))
    jobs = []
    objs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        return cmd, r

    with ThreadPoolExecutor(max_workers=8) as ex:
        for cmd, r in ex.map(run, jobs):
            if verbose or r.returncode != 0:
                sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
            if r.returncode != 0:
                raise RuntimeError("nvcc failed: " + cmd[-3])
    if force or jobs or _stale(LIB, objs):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                      "-ccbin", "/usr/bin/g++", "-lcudart", "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    build_host()
    return LIB


VDB_LIB = os.path.join(HERE, "libflipvdb.so")


def build_host():
    """Host-side pieces that need no GPU: the self-contained VDB-224 writer (used by the C++ driver
    and, through this small shared library, by the tests)."""
    src = os.path.join(HERE, "host", "vdb_writer.cc")
    hdr = os.path.join(HERE, "host", "vdb_writer.h")
    if _stale(VDB_LIB, [src, hdr]):
        cmd = ["/usr/bin/g++", "-std=c++11", "-O2", "-fPIC", "-shared", "-o", VDB_LIB, src, "-lz"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("vdb writer build failed")
    return VDB_LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
