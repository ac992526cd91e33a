"""examples/two_clusters.c: a consumer of the C ABI written in plain C.  It must compile and link
against libncdrift.so with a C compiler alone (CPU), and on a GPU print the decisions the oracle
makes for the same three caches."""
import os
import shutil
import subprocess

import pytest

from nexus_configuration_controller_b200 import _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build(tmp_path) -> str:
    cc = shutil.which("gcc") or shutil.which("cc")
    assert cc, "no C compiler in the image"
    libdir = os.path.dirname(_ffi.LIB_PATH)
    exe = str(tmp_path / "two_clusters")
    res = subprocess.run(
        [cc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
         os.path.join(ROOT, "examples", "two_clusters.c"), "-L", libdir, "-lncdrift", f"-Wl,-rpath,{libdir}", "-o", exe],
        capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    return exe


def test_c_example_compiles_and_links(tmp_path):
    _ffi.load()  # builds the library if it is stale
    exe = build(tmp_path)
    import torch

    if not torch.cuda.is_available():
        # without a GPU the program must fail loudly at ncd_init, not pretend
        res = subprocess.run([exe], capture_output=True, text=True, timeout=120)
        assert res.returncode == 2 and "ncd_init" in res.stderr


@pytest.mark.gpu
def test_c_example_decides_like_the_oracle(tmp_path):
    import oracle
    from nexus_configuration_controller_b200 import arena as A

    exe = build(tmp_path)
    res = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    O = A.CachedObject
    ctrl = A.pack([
        O(A.KIND_TEMPLATE, "nexus/algo", b"spec: image=algo:1.2 cpu=2", uid="uid-ctrl-algo",
          deps=[(A.KIND_SECRET, "nexus/algo-secret"), (A.KIND_CONFIGMAP, "nexus/algo-config")]),
        O(A.KIND_SECRET, "nexus/algo-secret", b"token=alpha", uid="uid-ctrl-secret"),
        O(A.KIND_CONFIGMAP, "nexus/algo-config", b"mode=fast", uid="uid-ctrl-config"),
    ])
    s0 = A.pack([
        O(A.KIND_CONFIGMAP, "nexus/algo-config", b"mode=fast", uid="uid-s0-config"),
        O(A.KIND_TEMPLATE, "nexus/algo", b"spec: image=algo:1.2 cpu=2", uid="uid-s0-algo"),
        O(A.KIND_SECRET, "nexus/algo-secret", b"token=bravo", uid="uid-s0-secret", owners=[(True, "uid-s0-algo")]),
    ])
    _, ents = oracle.scan(ctrl, [s0, A.pack([])])
    names = {1: "CREATE", 2: "UPDATE_DATA", 8: "ERR_ROGUE"}
    what = ["template nexus/algo", "secret nexus/algo-secret", "configmap nexus/algo-config"]
    lines = res.stdout.strip().splitlines()
    assert len(lines) == len(ents) + 1 and lines[-1].endswith("as expected")
    for line, (obj, dep, shard, code) in zip(lines, ents.tolist()):
        assert line.split() == f"shard {shard}: {what[dep]} {names[code]}".split()
