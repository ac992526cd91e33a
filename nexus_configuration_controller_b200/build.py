"""Builds libncdrift.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
from __future__ import annotations

import glob
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "lib", "libncdrift.so")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17", "-shared", "-Xcompiler", "-fPIC",
    "-diag-suppress", "177",
]


HOST_OUT = os.path.join(HERE, "lib", "libncdhost.so")


def host_sources():
    return sorted(glob.glob(os.path.join(CSRC, "host", "*.cpp")))


def build_host(force: bool = False) -> str:
    """libncdhost.so: the C++ host layer (canonicaliser, packer, DriftChecker, replay) over libncdrift.so."""
    deps = host_sources() + glob.glob(os.path.join(CSRC, "host", "*.hpp")) + glob.glob(os.path.join(HERE, "..", "include", "*.h")) + [OUT]
    if not force and os.path.exists(HOST_OUT) and all(os.path.getmtime(p) <= os.path.getmtime(HOST_OUT) for p in deps):
        return HOST_OUT
    cxx = shutil.which("g++") or "g++"
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-Wall", "-Wextra"] + host_sources() + [
        "-L" + os.path.dirname(OUT), "-lncdrift", "-Wl,-rpath,$ORIGIN", "-o", HOST_OUT]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("g++ failed:\n" + res.stdout + res.stderr)
    return HOST_OUT


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def stale() -> bool:
    if not os.path.exists(OUT):
        return True
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(HERE, "..", "include", "*.h"))
    return any(os.path.getmtime(p) > os.path.getmtime(OUT) for p in deps)


def build_native(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        build_host(False)
        return OUT
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    extra = os.environ.get("NCD_NVCC_EXTRA", "").split()  # experiment knobs, e.g. -DNCD_TMA_NPROD=6
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + sources() + ["-o", OUT]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    build_host(True)
    return OUT


if __name__ == "__main__":
    print(build_native(force=True, verbose=True))
