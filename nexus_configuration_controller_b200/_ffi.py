"""ctypes binding of libncdrift.so (include/ncdrift.h).

This is the same C ABI the Go controller binds through cgo; Python is only
the harness used by the tests and bench.py.  There is no fallback of any kind: if the shared
library is missing the import fails, and with no CUDA device ``ncd_init`` fails.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libncdrift.so")

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)


class ArenaView(C.Structure):
    """struct ncd_arena_view."""

    _fields_ = [
        ("nobj", C.c_uint64),
        ("payload_bytes", C.c_uint64),
        ("aux_bytes", C.c_uint64),
        ("nowner", C.c_uint64),
        ("ndep", C.c_uint64),
        ("payload", u8p),
        ("rec_off", u64p),
        ("rec_len", u32p),
        ("kind", u8p),
        ("key_off", u32p),
        ("key_len", u32p),
        ("uid_off", u32p),
        ("uid_len", u32p),
        ("own_start", u32p),
        ("own_is_template", u8p),
        ("own_uid_off", u32p),
        ("own_uid_len", u32p),
        ("dep_start", u32p),
        ("dep_kind", u8p),
        ("dep_key_off", u32p),
        ("dep_key_len", u32p),
        ("aux", u8p),
    ]


class Entry(C.Structure):
    """struct ncd_entry."""

    _fields_ = [("obj", C.c_uint32), ("dep", C.c_uint32), ("shard", C.c_uint32), ("code", C.c_uint32)]


class ScanStats(C.Structure):
    """struct ncd_scan_stats."""

    _fields_ = [
        ("join_ms", C.c_float),
        ("hash_ms", C.c_float),
        ("compare_ms", C.c_float),
        ("classify_ms", C.c_float),
        ("total_ms", C.c_float),
        ("rows", C.c_uint64),
        ("pairs", C.c_uint64),
        ("algo_bytes", C.c_uint64),
        ("ndiff", C.c_uint64),
        ("hash_collisions", C.c_uint64),
        ("launches", C.c_uint32),
        ("compare_variant", C.c_uint32),
    ]


class SynthCfg(C.Structure):
    """struct ncd_synth_cfg."""

    _fields_ = [
        ("groups", C.c_uint64),
        ("n_secrets", C.c_uint32),
        ("n_configmaps", C.c_uint32),
        ("tmpl_len_lo", C.c_uint32),
        ("tmpl_len_hi", C.c_uint32),
        ("dep_len_lo", C.c_uint32),
        ("dep_len_hi", C.c_uint32),
        ("seed", C.c_uint64),
        ("drift_seed", C.c_uint64),
        ("p_flip_ppm", C.c_uint32),
        ("p_missing_ppm", C.c_uint32),
        ("p_owner_ppm", C.c_uint32),
        ("p_rogue_ppm", C.c_uint32),
        ("p_len_ppm", C.c_uint32),
        ("part_mod", C.c_uint32),
        ("part_rem", C.c_uint32),
        ("shuffle", C.c_uint32),
    ]


# every symbol include/ncdrift.h declares: (restype, argtypes)
VP = C.c_void_p
PROTOTYPES = {
    "ncd_version": (C.c_char_p, []),
    "ncd_last_error": (C.c_char_p, []),
    "ncd_init": (C.c_int, [C.c_int, C.POINTER(VP)]),
    "ncd_destroy": (None, [VP]),
    "ncd_ctx_stream": (VP, [VP]),
    "ncd_ctx_device": (C.c_int, [VP]),
    "ncd_ctx_sync": (C.c_int, [VP]),
    "ncd_keyhash": (C.c_uint64, [C.c_uint32, C.c_char_p, C.c_uint32]),
    "ncd_arena_create": (C.c_int, [VP, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.POINTER(VP)]),
    "ncd_arena_destroy": (None, [VP]),
    "ncd_arena_host_view": (C.c_int, [VP, C.POINTER(ArenaView)]),
    "ncd_arena_commit": (C.c_int, [VP, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64]),
    "ncd_arenas_commit": (C.c_int, [C.c_int, C.POINTER(VP), u64p, u64p, u64p, u64p, u64p]),
    "ncd_arena_recommit": (C.c_int, [VP, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64]),
    "ncd_arena_sync_records": (C.c_int, [VP, C.c_uint64, u32p, C.c_uint64]),
    "ncd_arena_fetch_records": (C.c_int, [VP, C.c_uint64, u32p, VP, C.c_uint64, u64p]),
    "ncd_arena_compact": (C.c_int, [VP, u64p]),
    "ncd_arena_download": (C.c_int, [VP, C.POINTER(ArenaView)]),
    "ncd_arena_counts": (C.c_int, [VP, C.POINTER(ArenaView)]),
    "ncd_arena_rows": (C.c_int, [VP, u64p]),
    "ncd_scan": (C.c_int, [VP, VP, C.POINTER(VP), C.c_int, C.c_uint32, C.POINTER(VP)]),
    "ncd_scan_launch": (C.c_int, [VP, VP, C.POINTER(VP), C.c_int, C.c_uint32]),
    "ncd_scan_finish": (C.c_int, [VP, C.POINTER(VP)]),
    "ncd_scan_last_stats": (C.c_int, [VP, C.POINTER(ScanStats)]),
    "ncd_scan_stats_back": (C.c_int, [VP, C.c_uint64, C.POINTER(ScanStats)]),
    "ncd_scan_into": (C.c_int, [VP, VP, C.POINTER(VP), C.c_int, C.c_uint32, VP, C.c_uint64]),
    "ncd_scan_host": (C.c_int, [VP, VP, C.POINTER(VP), C.c_int, C.c_uint32, C.POINTER(C.POINTER(Entry)), u64p]),
    "ncd_find_unadopted": (C.c_int, [VP, VP, u32p, C.c_uint64, u64p]),
    "ncd_arenas_sync_records": (C.c_int, [C.c_int, C.POINTER(VP), u64p, C.POINTER(VP), u64p]),
    "ncd_arena_save": (C.c_int, [VP, C.c_char_p]),
    "ncd_arena_load": (C.c_int, [VP, C.c_char_p, C.c_uint32, C.c_uint64, C.POINTER(VP)]),
    "ncd_xchg_create": (C.c_int, [VP, C.c_int, C.c_int, C.c_uint64, C.POINTER(VP)]),
    "ncd_xchg_buffer": (C.c_int, [VP, C.POINTER(VP), u64p]),
    "ncd_xchg_handle": (C.c_int, [VP, VP]),
    "ncd_xchg_open": (C.c_int, [VP, VP]),
    "ncd_xchg_connect": (C.c_int, [VP, C.POINTER(VP)]),
    "ncd_scan_exchange": (C.c_int, [VP, VP, C.POINTER(VP), C.c_int, C.c_uint32, VP]),
    "ncd_xchg_wait": (C.c_int, [VP]),
    "ncd_xchg_result": (C.c_int, [VP, C.POINTER(VP), u64p]),
    "ncd_xchg_fetch": (C.c_int, [VP, VP, C.c_uint64]),
    "ncd_xchg_set_timeout": (C.c_int, [VP, C.c_double]),
    "ncd_xchg_destroy": (None, [VP]),
    "ncd_diff_view": (C.c_int, [VP, C.POINTER(C.POINTER(Entry)), u64p]),
    "ncd_diff_device": (C.c_int, [VP, C.POINTER(VP), u64p]),
    "ncd_diff_pack": (C.c_int, [VP, VP, C.c_uint64]),
    "ncd_diff_dirty_objects": (C.c_int, [VP, u32p, C.c_uint64, u64p]),
    "ncd_diff_codes": (C.c_int, [VP, u8p, C.c_uint64]),
    "ncd_diff_destroy": (None, [VP]),
    "ncd_find_orphans": (C.c_int, [VP, VP, C.POINTER(VP), C.c_int, u32p, C.c_uint64, u64p]),
    "ncd_hash_arena": (C.c_int, [VP, VP, C.c_uint32, VP, C.c_uint64, C.POINTER(C.c_float)]),
    "ncd_hash64_static": (C.c_int, [VP, VP, VP, C.c_uint64, VP, VP]),
    "ncd_synth_sizes": (C.c_int, [VP, C.POINTER(SynthCfg), C.c_int, u64p, u64p, u64p, u64p, u64p]),
    "ncd_synth_fill": (C.c_int, [VP, VP, C.POINTER(SynthCfg), C.c_int]),
}

_lib = None


def load() -> C.CDLL:
    """Loads libncdrift.so (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "There is no CPU or PyTorch fallback for the drift check."
            )
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)  # AttributeError here = header and library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class NcdError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"ncdrift status {status}: {message}")
        self.status = status


def check(status: int) -> None:
    if status != 0:
        raise NcdError(status, load().ncd_last_error().decode("utf-8", "replace"))
