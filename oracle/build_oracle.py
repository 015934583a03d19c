"""Build recipes for the test oracles (TEST INFRASTRUCTURE, not product code).

* ``build_cpu_oracle()``  -> oracle/_build/liboracle.so from oracle/raster_oracle.c (gcc).
* ``build_ref()``         -> oracle/_ref/libref_riggs.so: the reference's own CUDA translation units,
  compiled where they lie under /root/reference/submodules (never copied), linked with
  oracle/ref_shim.cu.  Skipped (returns None) when the reference tree is absent, e.g. on the GPU box,
  where the prebuilt file that travelled with the snapshot is used.

The reference's own build is a torch CUDAExtension (dgr/setup.py:17-34, knn/setup.py) that adds only
``-I third_party/glm``; we compile the same sources with plain nvcc defaults (device -O3, FMA
contraction on, no fast-math) for sm_100a and leave out its torch binding.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/submodules"
DGR = os.path.join(REF, "diff-gaussian-rasterization")
KNN = os.path.join(REF, "simple-knn")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_cpu_oracle(verbose=False, f64=False):
    """f64=True: the same source with -DORC_DOUBLE (every expression in double) -> liboracle_f64.so."""
    src = os.path.join(HERE, "raster_oracle.c")
    out_dir = os.path.join(HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "liboracle_f64.so" if f64 else "liboracle.so")
    if _newer(out, [src]):
        cmd = ["gcc", "-O2", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC", "-o", out, src, "-lm"]
        if f64:
            cmd.insert(1, "-DORC_DOUBLE")
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return out


def build_ref(verbose=False):
    out_dir = os.path.join(HERE, "_ref")
    out = os.path.join(out_dir, "libref_riggs.so")
    if not os.path.isdir(DGR):
        return out if os.path.exists(out) else None
    os.makedirs(out_dir, exist_ok=True)
    units = [
        (os.path.join(DGR, "cuda_rasterizer", "rasterizer_impl.cu"), "ref_rasterizer_impl.o"),
        (os.path.join(DGR, "cuda_rasterizer", "forward.cu"), "ref_forward.o"),
        (os.path.join(DGR, "cuda_rasterizer", "backward.cu"), "ref_backward.o"),
        (os.path.join(KNN, "simple_knn.cu"), "ref_simple_knn.o"),
        (os.path.join(HERE, "ref_shim.cu"), "ref_shim.o"),
    ]
    inc = ["-I", os.path.join(DGR, "third_party", "glm"), "-I", DGR, "-I", KNN]
    objs = []
    procs = []
    for src, o in units:
        obj = os.path.join(out_dir, o)
        objs.append(obj)
        if _newer(obj, [src]):
            cmd = [NVCC, "-std=c++17", *ARCH, "-Xcompiler", "-fPIC", "-w", *inc, "-c", src, "-o", obj]
            if verbose:
                print(" ".join(cmd))
            procs.append(subprocess.Popen(cmd))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("reference build failed")
    if _newer(out, objs):
        cmd = [NVCC, "-shared", *ARCH, "-o", out, *objs]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return out


if __name__ == "__main__":
    print(build_cpu_oracle(verbose=True))
    print(build_cpu_oracle(verbose=True, f64=True))
    print(build_ref(verbose=True))
