"""Builds lib/libmatcha_b200.so from csrc/*.cu with nvcc for sm_100a (in-tree, so the .so
travels with the repo snapshot).  Usage: python build.py [--force]"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-diag-suppress", "177"]


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
    # one nvcc -c per source, side by side (objects under build/, git-ignored), then one link
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    hdr_deps = glob.glob(os.path.join(HERE, "csrc", "*.cuh")) + glob.glob(os.path.join(HERE, "..", "include", "*.h")) + \
        [os.path.abspath(__file__)]
    jobs, objs = [], []
    for src in sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        fresh = os.path.exists(obj) and all(os.path.getmtime(d) <= os.path.getmtime(obj) for d in [src] + hdr_deps)
        if fresh and not force and not extra and not verbose:
            continue
        cmd = [NVCC] + FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        jobs.append((cmd, subprocess.Popen(cmd, cwd=HERE)))
    for cmd, p in jobs:
        if p.wait() != 0:
            raise subprocess.CalledProcessError(p.returncode, cmd)
    subprocess.run([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", lib_file()] + objs, check=True, cwd=HERE)
    return lib_file()


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
