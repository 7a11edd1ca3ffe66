"""In-tree build of the CUDA libraries (sm_100a only).

``librootsolve_b200.so``  the solver: C ABI of include/rootsolve_b200.h
``librs_models.so``       the bundled device models (BASELINE configs C1..C5)

Both are produced by plain ``nvcc`` invocations next to this file so that they
travel with the repository snapshot to the GPU box; nothing is JIT-compiled at
run time.  ``python -m rootsolve_b200.build`` rebuilds what is stale.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
INCLUDE = os.path.join(ROOT, "include")
CSRC = os.path.join(HERE, "csrc")
MODELS = os.path.join(HERE, "models")
OBJ = os.path.join(HERE, "_obj")

LIB_SOLVER = os.path.join(HERE, "librootsolve_b200.so")
LIB_MODELS = os.path.join(HERE, "librs_models.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ARCH + ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-Wno-deprecated-gpu-targets",
                 "-I", INCLUDE, "-I", CSRC, "-I", MODELS] + os.environ.get("RS_NVCC_DEFS", "").split()
# Translation units whose floating-point results are compared bit-for-bit with the
# oracle (residuals, FD quotients, error weights) must not contract a*b+c into FMA.
NO_FMA = {"fdjac.cu", "api.cu"}

SOLVER_SRCS = ["structure.cu", "jgroup.cu", "fdjac.cu", "krylov.cu", "mg.cu", "btd.cu", "api.cu", "comm.cu", "noalloc.cu"]
MODEL_SRCS = ["diff2d.cu", "diff3d.cu", "rd3d.cu", "rt1d.cu", "anox.cu", "diffmap.cu", "anoxforc.cu", "diff2db.cu", "o2bod.cu", "diff3dms.cu", "diff2dq.cu", "sediment.cu"]


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: the CUDA libraries cannot be built")
    return exe


def _nccl_flags():
    """Include/link flags for the NCCL that ships with torch (nvidia-nccl-cu12 wheel)."""
    try:
        import nvidia.nccl as pkg  # type: ignore
        base = os.path.dirname(pkg.__file__) if getattr(pkg, "__file__", None) else list(pkg.__path__)[0]
    except Exception:
        return None
    inc, lib = os.path.join(base, "include"), os.path.join(base, "lib")
    so = os.path.join(lib, "libnccl.so.2")
    if os.path.exists(os.path.join(inc, "nccl.h")) and os.path.exists(so):
        return inc, lib, so
    return None


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _headers():
    out = []
    for d in (INCLUDE, CSRC, MODELS):
        for f in os.listdir(d):
            if f.endswith((".h", ".cuh")):
                out.append(os.path.join(d, f))
    out.append(os.path.abspath(__file__))
    return out


def _run(cmd):
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if p.returncode != 0:
        raise RuntimeError("build failed:\n" + " ".join(cmd) + "\n" + p.stdout)
    return p.stdout


def _compile(src_dir, name, extra, verbose):
    os.makedirs(OBJ, exist_ok=True)
    src = os.path.join(src_dir, name)
    obj = os.path.join(OBJ, name.replace(".cu", ".o"))
    if _newer(obj, [src] + _headers()):
        out = _run([nvcc()] + COMMON + extra + ["-c", src, "-o", obj])
        if verbose and out.strip():
            print(out)
    return obj


def build_all(force: bool = False, verbose: bool = False):
    if force and os.path.isdir(OBJ):
        shutil.rmtree(OBJ)
    nccl = _nccl_flags()
    objs = []
    for s in SOLVER_SRCS:
        extra = ["-fmad=false"] if s in NO_FMA else []
        if s == "comm.cu" and nccl:
            extra += ["-DRS_WITH_NCCL", "-I", nccl[0]]
        objs.append(_compile(CSRC, s, extra, verbose))
    if force or _newer(LIB_SOLVER, objs):
        link = [nvcc(), "-shared", "-o", LIB_SOLVER] + objs + ARCH
        if nccl:
            # resolved at run time against the libnccl.so.2 torch has already loaded (or via rpath)
            link += ["-L", nccl[1], "-l:libnccl.so.2", "-Xlinker", "-rpath," + nccl[1]]
        _run(link)
    mobjs = [_compile(MODELS, s, ["-fmad=false"], verbose) for s in MODEL_SRCS]
    if force or _newer(LIB_MODELS, mobjs):
        _run([nvcc(), "-shared", "-o", LIB_MODELS] + mobjs + ARCH)
    return LIB_SOLVER, LIB_MODELS


if __name__ == "__main__":
    libs = build_all(force="--force" in sys.argv, verbose=True)
    print("\n".join(libs))
