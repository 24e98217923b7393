"""In-tree build of libdftd4.so (CUDA engine + C-API) for sm_100a.

    python -m dftd4_b200.build [--force]

nvcc cross-compiles without a GPU.  Outputs go to dftd4_b200/lib/ (git-ignored, shipped to the GPU box by gpurun).
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(LIBDIR, "obj")
LIB = os.path.join(LIBDIR, "libdftd4.so")

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
CXX = os.environ.get("CXX") or "g++"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# -fmad stays on (FMA contraction is wanted in the force kernels); the few operations that must round like the
# reference's un-contracted Fortran (cutoff tests) use __dmul_rn/__dadd_rn explicitly.
NVCC_FLAGS = ARCH + ["-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]
# host code that forms translation vectors / wrapped coordinates: no FMA contraction, like a stock gfortran build
CXX_FLAGS = ["-O2", "-ffp-contract=off", "-fPIC", "-std=c++17", "-Wall"]

CU_SOURCES = ["d4_engine.cu"]
CXX_SOURCES = ["d4_host.cpp", "d4_api.cpp"]


def _deps():
    out = []
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for f in os.listdir(root):
            if f.endswith((".cu", ".cuh", ".cpp", ".hpp", ".h", ".inc")):
                out.append(os.path.join(root, f))
    return out


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd, log=None):
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + res.stdout)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("build step failed: " + " ".join(cmd))
    return res.stdout


def build(force=False, verbose=False):
    os.makedirs(OBJDIR, exist_ok=True)
    deps = _deps()
    if not force and not _stale(LIB, deps):
        return LIB
    objs = []
    for src in CU_SOURCES:
        obj = os.path.join(OBJDIR, src + ".o")
        if force or _stale(obj, deps):
            _run([NVCC] + NVCC_FLAGS + ["-c", os.path.join(CSRC, src), "-o", obj], log=os.path.join(OBJDIR, src + ".ptxas.log"))
        objs.append(obj)
    cuda_inc = os.path.join(os.path.dirname(os.path.dirname(os.path.realpath(NVCC))), "include")
    for src in CXX_SOURCES:
        obj = os.path.join(OBJDIR, src + ".o")
        if force or _stale(obj, deps):
            _run([CXX] + CXX_FLAGS + ["-I", cuda_inc, "-c", os.path.join(CSRC, src), "-o", obj])
        objs.append(obj)
    _run([NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-ldl"])
    if verbose:
        print("built", LIB)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
