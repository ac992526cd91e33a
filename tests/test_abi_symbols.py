"""The C-ABI library loads without a GPU and exports every symbol include/ncdrift.h declares (CPU)."""
import ctypes as C
import os
import re

from nexus_configuration_controller_b200 import _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "ncdrift.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ncd_[a-z0-9_]+)\s*\(", src)))


def test_header_and_library_agree():
    names = declared_functions()
    assert len(names) >= 25
    lib = C.CDLL(_ffi.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in ncdrift.h but not exported: {missing}"
    # and the Python binding covers the same set
    assert sorted(_ffi.PROTOTYPES) == names


def test_struct_layouts_match_the_header():
    # sizes a C compiler gives the same declarations (all members naturally aligned)
    assert C.sizeof(_ffi.ArenaView) == 5 * 8 + 17 * 8
    assert C.sizeof(_ffi.Entry) == 16
    assert C.sizeof(_ffi.ScanStats) == 5 * 4 + 4 + 5 * 8 + 4 + 4  # 5 floats, pad, 5 u64, 2 u32
    assert C.sizeof(_ffi.SynthCfg) == 8 + 6 * 4 + 2 * 8 + 8 * 4


def test_no_gpu_means_a_loud_error_not_a_fallback():
    import torch

    if torch.cuda.is_available():
        return  # on the GPU box the context tests cover init
    lib = _ffi.load()
    h = C.c_void_p()
    rc = lib.ncd_init(0, C.byref(h))
    assert rc == -2 and not h.value
    assert b"no CUDA device" in lib.ncd_last_error()
    # host-only helper still works: the key hash is what the multi-GPU partition uses
    assert lib.ncd_keyhash(0, b"nexus/t-0000000001", 18) == lib.ncd_keyhash(0, b"nexus/t-0000000001", 18)
    assert lib.ncd_keyhash(0, b"nexus/t-0000000001", 18) != lib.ncd_keyhash(2, b"nexus/t-0000000001", 18)


def test_keyhash_matches_the_spec_used_for_partitioning():
    from tests import synth_spec

    lib = _ffi.load()
    for kind, key in [(0, b"nexus/t-0000000042"), (2, b"default/secret"), (3, b""), (1, b"a" * 300)]:
        assert lib.ncd_keyhash(kind, key, len(key)) == synth_spec.keyhash(kind, key)


def test_headers_are_plain_c_and_sizes_agree_with_a_c_compiler(tmp_path):
    """The boundary is a C ABI: both headers must compile as C99 on their own (no C++, no CUDA, no
    torch types), and the struct sizes a C compiler computes are the ones the ctypes mirror uses."""
    import shutil
    import subprocess

    cc = shutil.which("gcc") or shutil.which("cc")
    assert cc, "no C compiler in the image"
    src = tmp_path / "abi.c"
    src.write_text(
        '#include <stdio.h>\n#include "ncdrift.h"\n#include "ncdhost.h"\n'
        "int main(void) {\n"
        '  printf("%zu %zu %zu %zu %zu\\n", sizeof(ncd_arena_view), sizeof(ncd_entry), sizeof(ncd_scan_stats),\n'
        "         sizeof(ncd_synth_cfg), sizeof(ncd_orphan));\n"
        "  return 0;\n}\n"
    )
    exe = tmp_path / "abi"
    res = subprocess.run([cc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
                          str(src), "-o", str(exe)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    assert sizes[:4] == [C.sizeof(_ffi.ArenaView), C.sizeof(_ffi.Entry), C.sizeof(_ffi.ScanStats), C.sizeof(_ffi.SynthCfg)]
    assert sizes[4] == 8


def test_go_binding_and_docs_only_name_functions_the_header_declares():
    """go/ cannot be compiled here (no Go toolchain); at least every C function it calls, and every
    ncd_* function the integration guide mentions, must exist in the header with that name, and
    every constant the Go side reads must be defined."""
    declared = set(declared_functions())
    header = open(os.path.join(ROOT, "include", "ncdrift.h")).read()
    host_header = open(os.path.join(ROOT, "include", "ncdhost.h")).read()
    go_dir = os.path.join(ROOT, "go")
    used, consts = set(), set()
    for base, _, files in os.walk(go_dir):
        for f in files:
            if f.endswith(".go"):
                src = open(os.path.join(base, f)).read()
                used |= set(re.findall(r"\bC\.(ncd_[a-z0-9_]+)\s*\(", src))
                consts |= set(re.findall(r"\bC\.(NCD_[A-Z0-9_]+)\b", src))
    assert len(used) >= 20, "the Go binding went missing"
    assert not (used - declared), f"go/ calls functions the header does not declare: {sorted(used - declared)}"
    missing_consts = [c for c in consts if not re.search(rf"\b{c}\b", header)]
    assert not missing_consts, f"go/ reads constants the header does not define: {missing_consts}"
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    host_declared = set(re.findall(r"\b(ncdh_[a-z0-9_]+)\s*\(", re.sub(r"/\*.*?\*/", "", host_header, flags=re.S)))
    mentioned = set(re.findall(r"\b(ncdh?_[a-z0-9_]+)\s*\(", doc))
    unknown = {m for m in mentioned if m not in declared and m not in host_declared}
    assert not unknown, f"INTEGRATION.md mentions functions no header declares: {sorted(unknown)}"
