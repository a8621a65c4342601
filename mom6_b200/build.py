"""Build the CUDA shared library in-tree with nvcc for sm_100a.

    python -m mom6_b200.build [--force]

-fmad=false: no FMA contraction, so the device arithmetic is the reference's
operation-for-operation (SURVEY.md section 8a hazard 1).  Explicit __fma_rn is
used only inside the exact-division helpers of internal.cuh.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libmom6adv_b200.so")
SOURCES = ["kernels.cu", "plan.cu", "fast_x.cu", "fast_y.cu", "sparse.cu", "halo.cu", "obc.cu", "step.cu", "noalloc.cu", "hordiff.cu", "offline.cu"]
HEADERS = ["internal.cuh", "kernels.h", "fast_common.cuh", "obc.cuh", os.path.join("..", "..", "include", "mom6adv.h")]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-fmad=false", "-lineinfo",
              "-std=c++17", "-Xcompiler", "-fPIC", "-shared"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _nccl_flags():
    """Link against the NCCL that PyTorch ships (site-packages/nvidia/nccl), with an rpath to it: the host
    mirror and torch.distributed then share ONE libnccl.so.2 whichever is loaded first (the system 2.27 loaded
    first would leave torch's libtorch_cuda.so without its 2.28 symbols).  A pure Fortran host links the system
    library instead: falls back to -lnccl when the bundled copy is absent."""
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        base = list(spec.submodule_search_locations)[0] if spec and spec.submodule_search_locations else None
    except Exception:
        base = None
    if base and os.path.exists(os.path.join(base, "lib", "libnccl.so.2")) and not os.environ.get("MOM6ADV_SYSTEM_NCCL"):
        lib = os.path.join(base, "lib")
        return ["-I" + os.path.join(base, "include"), "-L" + lib, "-l:libnccl.so.2", "-Xlinker", "-rpath=" + lib]
    return ["-lnccl"]


def sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + [os.path.join(CSRC, h) for h in HEADERS]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return LIB
    extra = os.environ.get("MOM6ADV_NVCC_EXTRA", "").split()      # experiments: -DXMINB=5 ...
    cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + sources() + _nccl_flags()
    env = dict(os.environ)
    # the image exports CC/CXX pointing at a wrapper without libgomp; nvcc needs the system g++
    cmd += ["-ccbin", "/usr/bin/g++"] if os.path.exists("/usr/bin/g++") else []
    r = subprocess.run(cmd, cwd=CSRC, env=env, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
