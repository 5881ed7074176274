"""Builds libspatial_b200.so (sm_100a only) in-tree with nvcc. No torch dependency: plain CUDA runtime.

Every .cu is compiled to an object of its own (in parallel; objects are rebuilt only when the source, a header or the
flags changed) and the objects are linked into the one shared library the C-ABI lives in.
"""
import concurrent.futures
import glob
import hashlib
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
LIB = os.path.join(HERE, "libspatial_b200.so")

NVCC_FLAGS = [
    "-O3", "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-fmad=false",  # the reference is built without FMA contraction: products and sums round individually
    "-Xcompiler", "-fPIC,-fvisibility=hidden",
]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def headers():
    return sorted(glob.glob(os.path.join(CSRC, "*.cuh"))) + [os.path.join(HERE, "..", "include", "spatial_b200.h")]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(d) > t for d in sources() + headers())


def _compile(nvcc, flags, src, verbose):
    obj = os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")
    tag = obj + ".flags"
    want = hashlib.sha1(" ".join(flags).encode()).hexdigest()
    newest = max(os.path.getmtime(p) for p in [src] + headers())
    if (os.path.exists(obj) and os.path.getmtime(obj) > newest and os.path.exists(tag) and open(tag).read() == want):
        return obj
    cmd = [nvcc] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
    subprocess.check_call(cmd, cwd=CSRC)
    with open(tag, "w") as f:
        f.write(want)
    return obj


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("SB_NVCC_EXTRA", "").split()  # experiments only, e.g. -DBP_SLICE=131072u
    flags = NVCC_FLAGS + extra
    os.makedirs(OBJ, exist_ok=True)
    if force:
        for f in glob.glob(os.path.join(OBJ, "*.o")):
            os.remove(f)
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(lambda s: _compile(nvcc, flags, s, verbose), sources()))
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
