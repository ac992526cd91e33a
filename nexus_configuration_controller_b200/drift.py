"""Python harness over the C ABI: context, arenas, scans, diff lists.

Mirrors the Go-side wrapper of INTEGRATION.md (``DriftChecker``): ``Context.scan`` is what
``DriftChecker.Scan`` calls once per resync sweep, ``Diff.lookup`` what the patched call sites
of controller.go:539,600,725,795 consult instead of calling reflect.DeepEqual.  Everything that
computes lives in libncdrift.so; this file only moves pointers.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _ffi
from .arena import HostArena, MODE_DIRECT
from ._ffi import ArenaView, Entry, ScanStats, SynthCfg, check


class Context:
    """One GPU's drift-check context (ncd_ctx)."""

    def __init__(self, device: int = 0):
        self.lib = _ffi.load()
        h = C.c_void_p()
        check(self.lib.ncd_init(device, C.byref(h)))
        self.handle = h
        self.device = device

    def close(self) -> None:
        if self.handle:
            self.lib.ncd_destroy(self.handle)
            self.handle = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @property
    def stream(self) -> int:
        return int(self.lib.ncd_ctx_stream(self.handle) or 0)

    def sync(self) -> None:
        check(self.lib.ncd_ctx_sync(self.handle))

    # ---- arenas
    def arena_from_host(self, host: HostArena) -> "Arena":
        """Creates a pinned arena sized for `host`, copies it in and commits (H2D + index build)."""
        a = Arena(self, host.payload_bytes, host.nobj, host.aux_bytes, host.nowner, host.ndep, device_only=False)
        a.fill_from(host)
        a.commit(host.nobj, host.payload_bytes, host.aux_bytes, host.nowner, host.ndep)
        return a

    def synth_arena(self, cfg: SynthCfg, role: int) -> "Arena":
        """Device-only arena filled by the on-device generator (role -1 = controller, s = shard s)."""
        vals = [C.c_uint64() for _ in range(5)]
        check(self.lib.ncd_synth_sizes(self.handle, C.byref(cfg), role, *[C.byref(v) for v in vals]))
        payload, nobj, aux, nowner, ndep = [v.value for v in vals]
        a = Arena(self, payload, nobj, aux, nowner, ndep, device_only=True)
        check(self.lib.ncd_synth_fill(self.handle, a.handle, C.byref(cfg), role))
        return a

    # ---- scan
    def scan(self, ctrl: "Arena", shards: Sequence["Arena"], mode: int = MODE_DIRECT) -> "Diff":
        arr = (C.c_void_p * max(len(shards), 1))(*[s.handle for s in shards])
        out = C.c_void_p()
        check(self.lib.ncd_scan(self.handle, ctrl.handle, arr, len(shards), mode, C.byref(out)))
        return Diff(self, out, ctrl, len(shards))

    def scan_launch(self, ctrl: "Arena", shards: Sequence["Arena"], mode: int = MODE_DIRECT) -> None:
        arr = (C.c_void_p * max(len(shards), 1))(*[s.handle for s in shards])
        check(self.lib.ncd_scan_launch(self.handle, ctrl.handle, arr, len(shards), mode))
        self._pending = (ctrl, len(shards))

    def scan_finish(self) -> "Diff":
        out = C.c_void_p()
        check(self.lib.ncd_scan_finish(self.handle, C.byref(out)))
        ctrl, ns = self._pending
        return Diff(self, out, ctrl, ns)

    def scan_into(self, ctrl: "Arena", shards: Sequence["Arena"], mode: int, dev_ptr: int, cap_entries: int) -> None:
        """Asynchronous scan whose diff list lands in a caller-owned device buffer (header + entries)."""
        arr = (C.c_void_p * max(len(shards), 1))(*[s.handle for s in shards])
        check(self.lib.ncd_scan_into(self.handle, ctrl.handle, arr, len(shards), mode, C.c_void_p(dev_ptr), cap_entries))

    def sync_records_many(self, arenas: Sequence["Arena"], idx_lists: Sequence[np.ndarray], payload_bytes: Sequence[int]) -> None:
        """ncd_arenas_sync_records: value rewrites of several resident arenas, one round trip."""
        k = len(arenas)
        idx = [np.ascontiguousarray(i, dtype=np.uint32) for i in idx_lists]
        check(
            self.lib.ncd_arenas_sync_records(
                k,
                (C.c_void_p * max(k, 1))(*[a.handle for a in arenas]),
                (C.c_uint64 * max(k, 1))(*[len(i) for i in idx]),
                (C.c_void_p * max(k, 1))(*[i.ctypes.data for i in idx]),
                (C.c_uint64 * max(k, 1))(*[int(p) for p in payload_bytes]),
            )
        )

    def commit_many(self, arenas: Sequence["Arena"], counts: Sequence[Tuple[int, int, int, int, int]]) -> None:
        """ncd_arenas_commit: counts[k] = (nobj, payload_bytes, aux_bytes, nowner, ndep) of arenas[k]."""
        k = len(arenas)
        for a in arenas:
            a._meta = None
        cols = [(C.c_uint64 * max(k, 1))(*[int(c[j]) for c in counts]) for j in range(5)]
        check(self.lib.ncd_arenas_commit(k, (C.c_void_p * max(k, 1))(*[a.handle for a in arenas]), *cols))

    def stats_back(self, back: int) -> ScanStats:
        st = ScanStats()
        check(self.lib.ncd_scan_stats_back(self.handle, back, C.byref(st)))
        return st

    def find_orphans(self, ctrl: "Arena", shards: Sequence["Arena"], cap: int = 1 << 20) -> np.ndarray:
        """uint32[n, 2] of (shard, object index in that shard's arena) the controller does not hold."""
        arr = (C.c_void_p * max(len(shards), 1))(*[s.handle for s in shards])
        out = np.zeros((max(cap, 1), 2), dtype=np.uint32)
        n = C.c_uint64()
        check(self.lib.ncd_find_orphans(self.handle, ctrl.handle, arr, len(shards), out.ctypes.data_as(_ffi.u32p), cap, C.byref(n)))
        if n.value > cap:
            return self.find_orphans(ctrl, shards, cap=n.value)
        return out[: n.value].copy()

    def last_stats(self) -> ScanStats:
        st = ScanStats()
        check(self.lib.ncd_scan_last_stats(self.handle, C.byref(st)))
        return st

    def hash_arena(self, arena: "Arena", mode: int, fetch: bool = True) -> Tuple[Optional[np.ndarray], float]:
        """Digests of every record (uint64[n] for HASH64 and XXH64, uint8[n,32] for SHA256) and kernel ms."""
        n = arena.counts().nobj
        width = 32 if (mode & 0xFF) == 2 else 8
        out = np.zeros(max(n, 1) * width, dtype=np.uint8) if fetch else None
        ms = C.c_float()
        check(
            self.lib.ncd_hash_arena(
                self.handle,
                arena.handle,
                mode,
                out.ctypes.data_as(C.c_void_p) if fetch else None,
                out.nbytes if fetch else 0,
                C.byref(ms),
            )
        )
        if not fetch:
            return None, ms.value
        out = out[: n * width]
        return (out.view(np.uint64) if width == 8 else out.reshape(n, 32)), ms.value


class Arena:
    """ncd_arena: one cluster's cache on the GPU."""

    def __init__(self, ctx: Context, payload_cap, nobj_cap, aux_cap, nowner_cap, ndep_cap, device_only: bool):
        self.ctx = ctx
        h = C.c_void_p()
        check(
            ctx.lib.ncd_arena_create(
                ctx.handle, payload_cap, nobj_cap, aux_cap, nowner_cap, ndep_cap, 1 if device_only else 0, C.byref(h)
            )
        )
        self.handle = h

    def close(self) -> None:
        if self.handle:
            self.ctx.lib.ncd_arena_destroy(self.handle)
            self.handle = None

    def save(self, path: str) -> None:
        """Writes the committed arena to a snapshot file (ncd_arena_save)."""
        check(self.ctx.lib.ncd_arena_save(self.handle, os.fsencode(path)))

    @classmethod
    def load(cls, ctx: "Context", path: str, device_only: bool = True, payload_slack: int = 0) -> "Arena":
        """An arena from a snapshot file, checked, validated and indexed (ncd_arena_load)."""
        h = C.c_void_p()
        check(ctx.lib.ncd_arena_load(ctx.handle, os.fsencode(path), 1 if device_only else 0, payload_slack, C.byref(h)))
        a = cls.__new__(cls)
        a.ctx, a.handle = ctx, h
        return a

    def host_view(self) -> ArenaView:
        v = ArenaView()
        check(self.ctx.lib.ncd_arena_host_view(self.handle, C.byref(v)))
        return v

    def host_arrays(self) -> Dict[str, np.ndarray]:
        """numpy views over the pinned host arrays (capacity sized): what the packer writes into."""
        v = self.host_view()
        caps = {"payload": v.payload_bytes, "aux": v.aux_bytes, "own_start": v.nobj + 1, "dep_start": v.nobj + 1,
                "own_is_template": v.nowner, "own_uid_off": v.nowner, "own_uid_len": v.nowner,
                "dep_kind": v.ndep, "dep_key_off": v.ndep, "dep_key_len": v.ndep}
        out = {}
        for name, dt in HostArena.FIELDS:
            n = int(caps.get(name, v.nobj))
            out[name] = np.ctypeslib.as_array(getattr(v, name), shape=(max(n, 1),)) if n else np.zeros(0, dtype=dt)
        return out

    def fill_from(self, host: HostArena) -> None:
        """memcpy of a numpy arena image into the pinned host arrays (what the Go packer writes)."""
        v = self.host_view()
        for name, _ in HostArena.FIELDS:
            src = getattr(host, name)
            count = {
                "payload": host.payload_bytes,
                "aux": host.aux_bytes,
                "own_start": host.nobj + 1,
                "dep_start": host.nobj + 1,
                "own_is_template": host.nowner,
                "own_uid_off": host.nowner,
                "own_uid_len": host.nowner,
                "dep_kind": host.ndep,
                "dep_key_off": host.ndep,
                "dep_key_len": host.ndep,
            }.get(name, host.nobj)
            nbytes = count * src.itemsize
            if nbytes:
                C.memmove(getattr(v, name), src.ctypes.data, nbytes)

    def commit(self, nobj, payload_bytes, aux_bytes, nowner, ndep) -> None:
        self._meta = None
        check(self.ctx.lib.ncd_arena_commit(self.handle, nobj, payload_bytes, aux_bytes, nowner, ndep))

    def recommit(self, nobj, payload_bytes, aux_bytes, nowner, ndep, payload_dirty_from) -> None:
        """Structural change: re-upload the per-object arrays and only the appended payload."""
        self._meta = None
        check(self.ctx.lib.ncd_arena_recommit(self.handle, nobj, payload_bytes, aux_bytes, nowner, ndep, payload_dirty_from))

    def sync_records(self, obj_idx: np.ndarray, new_payload_bytes: int) -> None:
        """Pushes the records of `obj_idx` (rewritten in the pinned host arrays) to the device copy."""
        self._meta = None
        idx = np.ascontiguousarray(obj_idx, dtype=np.uint32)
        check(self.ctx.lib.ncd_arena_sync_records(self.handle, len(idx), idx.ctypes.data_as(_ffi.u32p), new_payload_bytes))

    def fetch_records(self, obj_idx) -> List[bytes]:
        """The canonical bytes of the named records, read back from the device (ncd_arena_fetch_records)."""
        idx = np.ascontiguousarray(obj_idx, dtype=np.uint32)
        n = len(idx)
        off = np.zeros(n + 1, dtype=np.uint64)
        lib = self.ctx.lib
        rc = lib.ncd_arena_fetch_records(self.handle, n, idx.ctypes.data_as(_ffi.u32p), None, 0, off.ctypes.data_as(_ffi.u64p))
        if rc not in (0, -4):
            check(rc)
        buf = np.zeros(max(int(off[n]), 16), dtype=np.uint8)
        check(lib.ncd_arena_fetch_records(self.handle, n, idx.ctypes.data_as(_ffi.u32p), buf.ctypes.data, buf.nbytes, off.ctypes.data_as(_ffi.u64p)))
        lens = self.download_meta().rec_len
        return [buf[int(off[k]) : int(off[k]) + int(lens[idx[k]])].tobytes() for k in range(n)]

    def download_meta(self) -> HostArena:
        """Every array of the arena except the payload (about 100 bytes per object)."""
        if getattr(self, "_meta", None) is None:
            c = self.counts()
            host = HostArena(c.nobj, 16, c.aux_bytes, c.nowner, c.ndep)
            host.payload_bytes = c.payload_bytes
            v = host.view()
            v.payload = None
            check(self.ctx.lib.ncd_arena_download(self.handle, C.byref(v)))
            self._meta = host
        return self._meta

    def compact(self) -> int:
        """Moves the live records to the front of the payload on both sides (ncd_arena_compact); returns the used bytes."""
        self._meta = None
        out = C.c_uint64()
        check(self.ctx.lib.ncd_arena_compact(self.handle, C.byref(out)))
        return out.value

    def counts(self) -> ArenaView:
        v = ArenaView()
        check(self.ctx.lib.ncd_arena_counts(self.handle, C.byref(v)))
        return v

    def rows(self) -> int:
        r = C.c_uint64()
        check(self.ctx.lib.ncd_arena_rows(self.handle, C.byref(r)))
        return r.value

    def download(self) -> HostArena:
        c = self.counts()
        host = HostArena(c.nobj, c.payload_bytes, c.aux_bytes, c.nowner, c.ndep)
        v = host.view()
        check(self.ctx.lib.ncd_arena_download(self.handle, C.byref(v)))
        return host


class Diff:
    """ncd_diff: the result of one scan."""

    def __init__(self, ctx: Context, handle, ctrl: Arena, nshards: int):
        self.ctx, self.handle, self.nshards = ctx, handle, nshards
        self.rows = ctrl.rows()
        self._codes: Optional[np.ndarray] = None

    def close(self) -> None:
        if self.handle:
            self.ctx.lib.ncd_diff_destroy(self.handle)
            self.handle = None

    def entries(self) -> np.ndarray:
        """The diff list as a uint32[n, 4] array of (obj, dep, shard, code), in (row, shard) order."""
        p = C.POINTER(Entry)()
        n = C.c_uint64()
        check(self.ctx.lib.ncd_diff_view(self.handle, C.byref(p), C.byref(n)))
        if n.value == 0:
            return np.zeros((0, 4), dtype=np.uint32)
        buf = C.cast(p, C.POINTER(C.c_uint32 * (4 * n.value))).contents
        return np.frombuffer(buf, dtype=np.uint32).reshape(n.value, 4).copy()

    def device_entries(self) -> Tuple[int, int]:
        p = C.c_void_p()
        n = C.c_uint64()
        check(self.ctx.lib.ncd_diff_device(self.handle, C.byref(p), C.byref(n)))
        return int(p.value or 0), n.value

    def pack_into(self, dev_ptr: int, cap_entries: int) -> None:
        """Header + entries into a caller-owned device buffer (the all-gather send buffer)."""
        check(self.ctx.lib.ncd_diff_pack(self.handle, C.c_void_p(dev_ptr), cap_entries))

    def dirty_objects(self) -> np.ndarray:
        """Controller indices of the templates / workgroups a sweep still has to reconcile."""
        n = C.c_uint64()
        check(self.ctx.lib.ncd_diff_dirty_objects(self.handle, None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 1), dtype=np.uint32)
        check(self.ctx.lib.ncd_diff_dirty_objects(self.handle, out.ctypes.data_as(_ffi.u32p), n.value, C.byref(n)))
        return out[: n.value]

    def codes(self) -> np.ndarray:
        """All decision codes, uint8[rows, nshards]."""
        if self._codes is None:
            n = self.rows * self.nshards
            out = np.zeros(max(n, 1), dtype=np.uint8)
            check(self.ctx.lib.ncd_diff_codes(self.handle, out.ctypes.data_as(_ffi.u8p), out.size))
            self._codes = out[:n].reshape(self.rows, self.nshards)
        return self._codes
