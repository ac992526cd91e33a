"""Builds the CPU test double of the host layer's backend (TEST INFRASTRUCTURE, never shipped):

  tests/_build/libncdhost_fake.so   csrc/host/*.cpp + fake_ncdrift.cpp + the oracle's decision code,
                                    same C ABI as libncdhost.so (include/ncdhost.h): host-logic tests on CPU
  tests/_build/host_stress[_tsan]   host_stress.cpp linked with the same sources, plain and with
                                    -fsanitize=thread
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(ROOT, "tests", "_build")
HOST = sorted(glob.glob(os.path.join(ROOT, "nexus_configuration_controller_b200", "csrc", "host", "*.cpp")))
COMMON = HOST + [os.path.join(HERE, "fake_ncdrift.cpp"), os.path.join(ROOT, "oracle", "ncd_oracle.cpp")]
DEPS = COMMON + glob.glob(os.path.join(ROOT, "nexus_configuration_controller_b200", "csrc", "host", "*.hpp")) + \
    glob.glob(os.path.join(ROOT, "include", "*.h")) + [os.path.join(HERE, "host_stress.cpp"), __file__]


def _stale(target: str) -> bool:
    return not os.path.exists(target) or any(os.path.getmtime(d) > os.path.getmtime(target) for d in DEPS)


def _run(cmd):
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("build failed: " + " ".join(cmd) + "\n" + res.stdout + res.stderr)


def build(tsan: bool = True) -> dict:
    os.makedirs(OUT, exist_ok=True)
    cxx = shutil.which("g++") or "g++"
    base = [cxx, "-std=c++17", "-pthread", "-Wall", "-Wextra", "-g"]
    out = {"lib": os.path.join(OUT, "libncdhost_fake.so"), "stress": os.path.join(OUT, "host_stress"),
           "stress_tsan": os.path.join(OUT, "host_stress_tsan")}
    if _stale(out["lib"]):
        _run(base + ["-O2", "-fPIC", "-shared"] + COMMON + ["-o", out["lib"]])
    if _stale(out["stress"]):
        _run(base + ["-O2"] + COMMON + [os.path.join(HERE, "host_stress.cpp"), "-o", out["stress"]])
    if tsan and _stale(out["stress_tsan"]):
        _run(base + ["-O1", "-fsanitize=thread"] + COMMON + [os.path.join(HERE, "host_stress.cpp"), "-o", out["stress_tsan"]])
    return out


if __name__ == "__main__":
    print(build())
