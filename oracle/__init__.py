"""CPU oracle of the drift-check hot path — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs may
import this package.  Nothing under nexus_configuration_controller_b200/ does: the product path
has no CPU route.  Sources: ncd_oracle.cpp (arena-level decisions), go_values.cpp (Go value
model, reflect.DeepEqual, canonical encoder), ncd_hashes.cpp (digests); built by
``make -C oracle`` into oracle/_build/libncd_oracle.so.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "libncd_oracle.so")
_lib = None


def build() -> None:
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        from nexus_configuration_controller_b200._ffi import ArenaView, Entry  # struct layouts only

        lib = C.CDLL(LIB_PATH)
        avp = C.POINTER(ArenaView)
        lib.oracle_rows.restype = C.c_uint64
        lib.oracle_rows.argtypes = [avp]
        lib.oracle_scan.restype = C.c_uint64
        lib.oracle_scan.argtypes = [avp, C.POINTER(avp), C.c_int, C.POINTER(C.c_uint8), C.POINTER(Entry), C.c_int, C.POINTER(C.c_uint64)]
        lib.oracle_scan_timed.restype = C.c_double
        lib.oracle_scan_timed.argtypes = [avp, C.POINTER(avp), C.c_int, C.POINTER(C.c_uint8), C.c_int, C.c_int, C.POINTER(C.c_uint64)]
        lib.oracle_unadopted.restype = C.c_uint64
        lib.oracle_unadopted.argtypes = [avp, C.POINTER(C.c_uint32), C.c_uint64]
        lib.oracle_fleet_create.restype = C.c_void_p
        lib.oracle_fleet_create.argtypes = [C.c_int]
        lib.oracle_fleet_destroy.restype = None
        lib.oracle_fleet_destroy.argtypes = [C.c_void_p]
        lib.oracle_fleet_apply.restype = C.c_int
        lib.oracle_fleet_apply.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int]
        lib.oracle_fleet_sweep.restype = C.c_double
        lib.oracle_fleet_sweep.argtypes = [C.c_void_p, C.c_int] + [C.POINTER(C.c_uint64)] * 4
        lib.oracle_h64.restype = C.c_uint64
        lib.oracle_h64.argtypes = [C.c_char_p, C.c_uint64]
        lib.oracle_sha256.restype = None
        lib.oracle_sha256.argtypes = [C.c_char_p, C.c_uint64, C.POINTER(C.c_uint8)]
        lib.oracle_xxh64.restype = C.c_uint64
        lib.oracle_xxh64.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64]
        lib.oracle_h64_arena.restype = None
        lib.oracle_h64_arena.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        lib.oracle_sha256_arena.restype = None
        lib.oracle_sha256_arena.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        lib.oracle_values_error.restype = C.c_char_p
        lib.oracle_deep_equal_bench.restype = C.c_double
        lib.oracle_deep_equal_bench.argtypes = [C.c_char_p, C.c_uint64, C.c_char_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(C.c_int)]
        lib.oracle_deep_equal.restype = C.c_int
        lib.oracle_deep_equal.argtypes = [C.c_char_p, C.c_uint64, C.c_char_p, C.c_uint64]
        lib.oracle_canon.restype = C.c_int64
        lib.oracle_canon.argtypes = [C.c_char_p, C.c_uint64, C.c_char_p, C.c_uint64]
        _lib = lib
    return _lib


def deep_equal_timed(a_wire: bytes, b_wire: bytes, npairs: int, reps: int, nthreads: int) -> Tuple[float, bool]:
    """Seconds `nthreads` threads take for `reps` sweeps of DeepEqual over `npairs` separately
    allocated (already parsed) copies of the two values; and whether every comparison said equal."""
    eq = C.c_int()
    sec = load().oracle_deep_equal_bench(a_wire, len(a_wire), b_wire, len(b_wire), npairs, reps, nthreads, C.byref(eq))
    if sec < 0:
        raise ValueError(load().oracle_values_error().decode())
    return sec, bool(eq.value)


def deep_equal(a_wire: bytes, b_wire: bytes) -> bool:
    """reflect.DeepEqual of two Go values given in the transport form of tests/govalues.py."""
    rc = load().oracle_deep_equal(a_wire, len(a_wire), b_wire, len(b_wire))
    if rc < 0:
        raise ValueError(load().oracle_values_error().decode())
    return rc == 1


def canon(wire: bytes) -> bytes:
    """Canonical encoding of a Go value (oracle's own statement of the encoder)."""
    lib = load()
    n = lib.oracle_canon(wire, len(wire), None, 0)
    if n < 0:
        raise ValueError(lib.oracle_values_error().decode())
    buf = C.create_string_buffer(max(n, 1))
    lib.oracle_canon(wire, len(wire), buf, n)
    return buf.raw[:n]


def _views(ctrl, shards):
    from nexus_configuration_controller_b200._ffi import ArenaView

    cv = ctrl.view()
    svs = [s.view() for s in shards]
    arr = (C.POINTER(ArenaView) * max(len(svs), 1))(*[C.pointer(v) for v in svs])
    return cv, svs, arr


def scan(ctrl, shards: Sequence, nthreads: int = 1) -> Tuple[np.ndarray, np.ndarray]:
    """Decision codes uint8[rows, nshards] and the diff list uint32[n, 4] for host arenas."""
    from nexus_configuration_controller_b200._ffi import Entry

    lib = load()
    cv, svs, arr = _views(ctrl, shards)
    rows = lib.oracle_rows(C.byref(cv))
    S = len(shards)
    codes = np.zeros(max(rows * S, 1), dtype=np.uint8)
    ents = (Entry * max(rows * S, 1))()
    n = lib.oracle_scan(C.byref(cv), arr, S, codes.ctypes.data_as(C.POINTER(C.c_uint8)), ents, nthreads, None)
    e = np.frombuffer(ents, dtype=np.uint32).reshape(-1, 4)[:n].copy()
    return codes[: rows * S].reshape(rows, S), e


def scan_timed(ctrl, shards: Sequence, nthreads: int, reps: int) -> Tuple[float, np.ndarray, int]:
    """(seconds per sweep, codes, controller-side bytes compared per sweep)."""
    lib = load()
    cv, svs, arr = _views(ctrl, shards)
    rows = lib.oracle_rows(C.byref(cv))
    S = len(shards)
    codes = np.zeros(max(rows * S, 1), dtype=np.uint8)
    nbytes = C.c_uint64()
    sec = lib.oracle_scan_timed(C.byref(cv), arr, S, codes.ctypes.data_as(C.POINTER(C.c_uint8)), nthreads, reps, C.byref(nbytes))
    return sec, codes[: rows * S].reshape(rows, S), nbytes.value


def h64(data: bytes) -> int:
    return load().oracle_h64(data, len(data))


def sha256(data: bytes) -> bytes:
    out = (C.c_uint8 * 32)()
    load().oracle_sha256(data, len(data), out)
    return bytes(out)


def xxh64(data: bytes, seed: int = 0) -> int:
    return load().oracle_xxh64(data, len(data), seed)


def h64_arena(a) -> np.ndarray:
    out = np.zeros(max(a.nobj, 1), dtype=np.uint64)
    load().oracle_h64_arena(a.payload.ctypes.data, a.rec_off.ctypes.data, a.rec_len.ctypes.data, a.nobj, out.ctypes.data)
    return out[: a.nobj]


def sha256_arena(a) -> np.ndarray:
    out = np.zeros(max(a.nobj, 1) * 32, dtype=np.uint8)
    load().oracle_sha256_arena(a.payload.ctypes.data, a.rec_off.ctypes.data, a.rec_len.ctypes.data, a.nobj, out.ctypes.data)
    return out[: a.nobj * 32].reshape(a.nobj, 32)


def unadopted(ctrl) -> np.ndarray:
    """uint32[n, 2] of (template index, dep) whose controller-side dependent lacks the template's ownerReference."""
    lib = load()
    cv = ctrl.view()
    n = lib.oracle_unadopted(C.byref(cv), None, 0)
    out = np.zeros((max(n, 1), 2), dtype=np.uint32)
    lib.oracle_unadopted(C.byref(cv), out.ctypes.data_as(C.POINTER(C.c_uint32)), n)
    return out[:n]


class Fleet:
    """Every cluster's cache as parsed objects + the handlers' sweep over them (oracle/go_values.cpp "fleet"):
    the CPU arm of the steady-state bench.  Events are ncdh_event arrays (include/ncdhost.h)."""

    def __init__(self, nshards: int):
        self.lib = load()
        self.h = C.c_void_p(self.lib.oracle_fleet_create(nshards))

    def close(self):
        if self.h:
            self.lib.oracle_fleet_destroy(self.h)
            self.h = None

    def apply(self, events_ptr, n: int, threads: int = 1) -> None:
        if self.lib.oracle_fleet_apply(self.h, C.cast(events_ptr, C.c_void_p), n, threads) != 0:
            raise ValueError(self.lib.oracle_values_error().decode())

    def sweep(self, threads: int = 1):
        """(seconds, decisions, diff entries, dirty objects, diff-list checksum)."""
        v = [C.c_uint64() for _ in range(4)]
        sec = self.lib.oracle_fleet_sweep(self.h, threads, *[C.byref(x) for x in v])
        return (sec,) + tuple(x.value for x in v)
