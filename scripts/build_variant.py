"""Builds a VARIANT of libspatial_b200.so for A/B measurements: the regular objects of duckdb-spatial_b200/_obj plus the
named .cu files recompiled with extra flags, linked to duckdb-spatial_b200/_variants/<name>.so (git-ignored *.so; it
travels to the GPU box). Use with SPATIAL_B200_LIB=<path>.

    python scripts/build_variant.py NAME "-DOS_MIN_CTAS=4 -DOS_LATE_VALS=2" radix_sort.cu [more.cu ...]
"""
import importlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
B = importlib.import_module("duckdb-spatial_b200.build")


def main():
    name, extra, files = sys.argv[1], sys.argv[2].split(), sys.argv[3:]
    B.build()
    out_dir = os.path.join(B.HERE, "_variants")
    os.makedirs(out_dir, exist_ok=True)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    for src in B.sources():
        base = os.path.basename(src)
        if base in files:
            obj = os.path.join(out_dir, name + "_" + base[:-3] + ".o")
            subprocess.check_call([nvcc] + B.NVCC_FLAGS + extra + ["-c", src, "-o", obj], cwd=B.CSRC)
        else:
            obj = os.path.join(B.OBJ, base[:-3] + ".o")
        objs.append(obj)
    lib = os.path.join(out_dir, name + ".so")
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib] + objs)
    print(lib)


if __name__ == "__main__":
    main()
