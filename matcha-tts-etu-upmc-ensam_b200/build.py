"""Builds lib/libmatcha_b200.so from csrc/*.cu with nvcc for sm_100a (in-tree, so the .so
travels with the repo snapshot).  Usage: python build.py [--force]"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-shared", "-Xcompiler", "-fPIC", "-diag-suppress", "177"]


def sources():
    return sorted(glob.glob(os.path.join(HERE, "csrc", "*.cu")))


def lib_file():
    return os.path.join(HERE, "lib", "libmatcha_b200.so")


def up_to_date():
    out = lib_file()
    if not os.path.exists(out):
        return False
    deps = sources() + glob.glob(os.path.join(HERE, "csrc", "*.cuh")) + \
        glob.glob(os.path.join(HERE, "..", "include", "*.h")) + [os.path.abspath(__file__)]
    return all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps)


def build(force=False, verbose=False):
    if not force and up_to_date():
        return lib_file()
    os.makedirs(os.path.join(HERE, "lib"), exist_ok=True)
    extra = os.environ.get("MB200_NVCC_EXTRA", "").split()          # e.g. -DSOME_PROFILING_SWITCH=1
    cmd = [NVCC] + FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", lib_file()] + sources()
    subprocess.run(cmd, check=True, cwd=HERE)
    return lib_file()


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
