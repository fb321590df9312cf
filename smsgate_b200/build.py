"""In-tree builds: the CUDA library (nvcc, sm_100a only) and the two test-only
CPU artefacts (the host check of the device functions, the C part of nothing
else).  `python -m smsgate_b200.build` builds everything that is stale."""
from __future__ import annotations

import os
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libsmsgate_b200.so")
HOSTCHECK_SRC = os.path.join(ROOT, "tests", "hostcheck", "hostcheck.cpp")
HOSTCHECK_LIB = os.path.join(ROOT, "tests", "hostcheck", "libhostcheck.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared", "--use_fast_math", "-Xptxas", "-v"]


def _stale(target: str, sources: list) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _csrc_files() -> list:
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh", ".h"))] + \
           [os.path.join(ROOT, "include", "smsgate_b200.h")]


def build_variant(name: str, defines: list) -> str:
    """Kernel experiments: the same library compiled with extra -D flags into smsgate_b200/variants/lib<name>.so;
    SMSGATE_B200_LIB=<path> makes _lib load it (scripts/gpu_variants.sh times them side by side)."""
    out_dir = os.path.join(PKG, "variants")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, f"lib{name}.so")
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + [f"-D{d}" for d in defines] + ["-o", out, os.path.join(CSRC, "smsgate_b200.cu")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(f"nvcc failed building variant {name}")
    with open(os.path.join(out_dir, f"{name}.ptxas.txt"), "w") as f:
        f.write(res.stderr)
    return out


def build_cuda(force: bool = False, verbose: bool = False) -> str:
    srcs = _csrc_files()
    if force or _stale(LIB, srcs):
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        cmd = [nvcc] + NVCC_FLAGS + ["-o", LIB, os.path.join(CSRC, "smsgate_b200.cu")]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or res.returncode:
            sys.stderr.write(res.stdout + res.stderr)
        if res.returncode:
            raise RuntimeError("nvcc failed building libsmsgate_b200.so")
        with open(os.path.join(PKG, "ptxas_info.txt"), "w") as f:
            f.write(res.stderr)
    return LIB


def build_hostcheck(force: bool = False) -> str:
    """TEST ONLY: the __host__ __device__ cores compiled for the CPU with g++,
    so the `-m "not gpu"` suite can check the device logic without a GPU."""
    srcs = _csrc_files() + [HOSTCHECK_SRC]
    if force or _stale(HOSTCHECK_LIB, srcs):
        cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", HOSTCHECK_LIB, HOSTCHECK_SRC,
               "-I", CSRC]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("g++ failed building libhostcheck.so")
    return HOSTCHECK_LIB


if __name__ == "__main__":
    print(build_cuda(force="--force" in sys.argv, verbose=True))
    print(build_hostcheck(force="--force" in sys.argv))
