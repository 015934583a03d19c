"""Builds riggs_b200/lib/libriggs_b200.so (sm_100a) with nvcc, in-tree.

The library has no torch dependency: it is a plain CUDA shared object with the C ABI of
include/riggs_b200.h, so the same file serves Python (ctypes), C and C++ hosts.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(LIBDIR, "obj")
LIB = os.path.join(LIBDIR, "libriggs_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-O3", "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")]
    hs.append(os.path.join(os.path.dirname(HERE), "include", "riggs_b200.h"))
    return hs


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(verbose: bool = False, force: bool = False, variant: str = "", defines=()) -> str:
    """variant / defines: a tuning build -- the same sources with extra -D flags, linked to lib/libriggs_b200_<variant>.so
    (selected at run time with RG_LIB=<path>); the product library is the plain build."""
    objdir = OBJDIR + ("_" + variant if variant else "")
    lib = LIB if not variant else os.path.join(LIBDIR, f"libriggs_b200_{variant}.so")
    flags = FLAGS + [f"-D{d}" for d in defines]
    os.makedirs(objdir, exist_ok=True)
    hdrs = _headers()
    jobs = []
    objs = []
    for src in _sources():
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src[:-3] + ".o")
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            jobs.append((s, o))

    def run(job):
        s, o = job
        cmd = [NVCC, *flags, "-c", s, "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = o[:-2] + ".ptxas.log"
        with open(log, "w") as f:
            f.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {s}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print("compiled", os.path.basename(s))
        return o

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(run, jobs))
    if force or jobs or _stale(lib, objs):
        cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib, *objs]
        subprocess.check_call(cmd)
        if verbose:
            print("linked", lib)
    return lib


if __name__ == "__main__":
    # python build.py [--force] [--variant NAME -DX=1 -DY=2 ...]
    args = sys.argv[1:]
    variant = args[args.index("--variant") + 1] if "--variant" in args else ""
    defines = [a[2:] for a in args if a.startswith("-D")]
    print(build(verbose=True, force="--force" in args, variant=variant, defines=defines))
