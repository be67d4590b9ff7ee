#!/usr/bin/env python
"""Builds libb200dp.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python build.py [--force] [--verbose]

Output: neural-networks-parallel-training-with-mpi_b200/lib/libb200dp.so (git-ignored; travels to
the GPU box with the gpurun snapshot).  Objects are cached under build/ and rebuilt when a source
or header is newer."""
import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "neural-networks-parallel-training-with-mpi_b200")
OUT = os.path.join(PKG, "lib", "libb200dp.so")
OBJ = os.path.join(ROOT, "build", "obj")
NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v",
         "-I", os.path.join(ROOT, "include")]


EXTRA = {}      # per-file nvcc flags (none at present)
FLAGS += os.environ.get("B200DP_NVCC_EXTRA", "").split()      # A/B builds: -D switches (with --force)


def sources():
    return sorted(glob.glob(os.path.join(ROOT, "csrc", "*.cu")))


def stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    headers = glob.glob(os.path.join(ROOT, "csrc", "*.cuh")) + glob.glob(os.path.join(ROOT, "include", "*.h"))
    jobs = []
    for src in sources():
        obj = os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")
        if force or stale(obj, [src, *headers]):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        r = subprocess.run([NVCC, *FLAGS, *EXTRA.get(os.path.basename(src), []), "-c", src, "-o", obj],
                           capture_output=True, text=True)
        return src, r

    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for src, r in ex.map(compile_one, jobs):
                log = os.path.join(OBJ, os.path.basename(src)[:-3] + ".ptxas.log")
                with open(log, "w") as f:
                    f.write(r.stderr)
                if verbose or r.returncode != 0:
                    sys.stderr.write(r.stdout + r.stderr)
                if r.returncode != 0:
                    raise RuntimeError(f"nvcc failed on {src}")
    objs = [os.path.join(OBJ, os.path.basename(s)[:-3] + ".o") for s in sources()]
    if force or jobs or stale(OUT, objs):
        r = subprocess.run([NVCC, "-shared", "-o", OUT, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
                            "-lcudart_static", "-lpthread", "-ldl", "-lrt"], capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    out = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(out)
