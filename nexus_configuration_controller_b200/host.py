"""ctypes binding of libncdhost.so (include/ncdhost.h): the host layer the Go controller would own.

``DriftChecker`` mirrors the Go interface of INTEGRATION.md section 2: fill the cluster images the
way the reference's fixture fills the informer indexers (controller_test.go:546-576), ``scan()``,
then ``lookup()`` / ``plan_template()`` / ``plan_workgroup()``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

from . import _ffi

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libncdhost.so")
VP = C.c_void_p
CPP = C.POINTER(C.c_char_p)
NOT_FOUND, STALE = 1, 2  # NCDH_NOT_FOUND, NCDH_STALE
EV_PUT, EV_DELETE = 0, 1


class Event(C.Structure):
    """struct ncdh_event."""

    _fields_ = [
        ("cluster", C.c_int32), ("kind", C.c_uint32), ("op", C.c_uint32),
        ("nowners", C.c_int32), ("nsecrets", C.c_int32), ("nconfigmaps", C.c_int32),
        ("ns", C.c_char_p), ("name", C.c_char_p), ("uid", C.c_char_p),
        ("value", C.c_void_p), ("value_len", C.c_uint64),
        ("owner_kinds", CPP), ("owner_uids", CPP), ("secret_names", CPP), ("configmap_names", CPP),
    ]


class ScanTiming(C.Structure):
    """struct ncdh_scan_timing."""

    _fields_ = [(n, C.c_double) for n in ("canon_ms", "pack_ms", "upload_ms", "scan_ms", "publish_ms", "total_ms", "gpu_scan_ms")] + [
        (n, C.c_uint64) for n in ("h2d_bytes", "d2h_bytes", "records_synced", "full_uploads", "structural_uploads", "events", "compactions", "records_in_place")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


PROTOTYPES = {
    "ncdh_last_error": (C.c_char_p, []),
    "ncdh_canon": (C.c_int64, [C.c_char_p, C.c_uint64, C.c_char_p, C.c_uint64]),
    "ncdh_checker_create": (C.c_int, [C.c_int, C.c_uint32, C.c_int, C.POINTER(VP)]),
    "ncdh_checker_destroy": (None, [VP]),
    "ncdh_cluster_add": (C.c_int, [VP, C.c_int, C.c_uint32, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint64,
                                   CPP, CPP, C.c_int, CPP, C.c_int, CPP, C.c_int]),
    "ncdh_cluster_remove": (C.c_int, [VP, C.c_int, C.c_uint32, C.c_char_p, C.c_char_p]),
    "ncdh_apply_events": (C.c_int, [VP, C.POINTER(Event), C.c_uint64, C.c_int]),
    "ncdh_plan_adoption": (C.c_int64, [VP, C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint64]),
    "ncdh_last_timing": (C.c_int, [VP, C.POINTER(ScanTiming)]),
    "ncdh_batch_from_arrays": (C.c_int, [C.POINTER(_ffi.ArenaView), C.POINTER(_ffi.ArenaView), C.c_int, _ffi.u32p, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(VP)]),
    "ncdh_batch_events": (C.POINTER(Event), [VP, C.POINTER(C.c_uint64)]),
    "ncdh_batch_value_bytes": (C.c_uint64, [VP]),
    "ncdh_batch_free": (None, [VP]),
    "ncdh_diff_digest": (C.c_int, [VP, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "ncdh_scan": (C.c_int, [VP]),
    "ncdh_lookup": (C.c_int, [VP, C.c_uint32, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_uint32)]),
    "ncdh_plan_template": (C.c_int64, [VP, C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint64]),
    "ncdh_plan_workgroup": (C.c_int64, [VP, C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint64]),
    "ncdh_sweep_queue": (C.c_int64, [VP, C.c_char_p, C.c_uint64]),
    "ncdh_upload_counts": (C.c_int, [VP, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "ncdh_last_counts": (C.c_int, [VP, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
}
_lib = None


def bind(lib: C.CDLL) -> C.CDLL:
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        _ffi.load()  # libncdrift.so first (same directory, also found through the rpath)
        _lib = bind(C.CDLL(LIB_PATH))
    return _lib


class HostError(RuntimeError):
    pass


class Stale(HostError):
    """The caches changed after the snapshot's scan began (NCDH_STALE): scan again, then ask again."""


def _err(lib=None) -> HostError:
    return HostError((lib or load()).ncdh_last_error().decode("utf-8", "replace"))


def canon(value_wire: bytes) -> bytes:
    """Canonical bytes of one Go value (transport form in): what the packer appends to an arena."""
    lib = load()
    n = lib.ncdh_canon(value_wire, len(value_wire), None, 0)
    if n < 0:
        raise _err()
    buf = C.create_string_buffer(max(n, 1))
    lib.ncdh_canon(value_wire, len(value_wire), buf, n)
    return buf.raw[:n]


def _strs(items: Sequence[str]):
    return (C.c_char_p * max(len(items), 1))(*[s.encode() for s in items])


def make_event(cluster: int, kind: int, namespace: str, name: str, uid: str = "", value_wire: bytes = b"",
               owner_refs: Sequence[Tuple[str, str]] = (), secret_names: Sequence[str] = (),
               configmap_names: Sequence[str] = (), op: int = EV_PUT, keep: Optional[list] = None) -> Event:
    """One ncdh_event; the buffers it points into are appended to `keep` (hold it until the call returns)."""
    e = Event()
    bufs = [namespace.encode(), name.encode(), uid.encode(), bytes(value_wire),
            _strs([k for k, _ in owner_refs]), _strs([u for _, u in owner_refs]), _strs(list(secret_names)), _strs(list(configmap_names))]
    e.cluster, e.kind, e.op = cluster, kind, op
    e.nowners, e.nsecrets, e.nconfigmaps = len(owner_refs), len(secret_names), len(configmap_names)
    e.ns, e.name, e.uid = bufs[0], bufs[1], bufs[2]
    e.value, e.value_len = C.cast(C.c_char_p(bufs[3]), C.c_void_p), len(bufs[3])
    e.owner_kinds, e.owner_uids = C.cast(bufs[4], CPP), C.cast(bufs[5], CPP)
    e.secret_names, e.configmap_names = C.cast(bufs[6], CPP), C.cast(bufs[7], CPP)
    if keep is not None:
        keep.append(bufs)
    return e


class EventBatch:
    """ncdh_event_batch: synthetic informer events made from arena-shaped host arrays (bench / tests)."""

    def __init__(self, meta, values=None, cluster: int = -1, obj_idx=None, salt: int = 0, threads: int = 0, lib=None):
        """meta / values: HostArena (arena.py); obj_idx: uint32 indices or None for every object."""
        import numpy as np

        self.lib = lib if lib is not None else load()
        mv = meta.view()
        vv = values.view() if values is not None else None
        idx = None if obj_idx is None else np.ascontiguousarray(obj_idx, dtype=np.uint32)
        h = VP()
        rc = self.lib.ncdh_batch_from_arrays(C.byref(mv), C.byref(vv) if vv is not None else None, cluster,
                                             idx.ctypes.data_as(_ffi.u32p) if idx is not None else None,
                                             0 if idx is None else len(idx), salt, threads, C.byref(h))
        if rc != 0:
            raise HostError("ncdh_batch_from_arrays: bad arrays or index")
        self.h = h
        n = C.c_uint64()
        self.events = self.lib.ncdh_batch_events(self.h, C.byref(n))
        self.n = n.value
        self.value_bytes = self.lib.ncdh_batch_value_bytes(self.h)

    def close(self):
        if self.h:
            self.lib.ncdh_batch_free(self.h)
            self.h = None

    def __del__(self):
        self.close()


class DriftChecker:
    def __init__(self, nshards: int, device: int = 0, mode: int = 0, lib: Optional[C.CDLL] = None):
        """lib: another build of the same ABI (the CPU test double of tests/fake_backend); default libncdhost.so."""
        self.lib = lib if lib is not None else load()
        h = VP()
        if self.lib.ncdh_checker_create(device, mode, nshards, C.byref(h)) != 0:
            raise _err(self.lib)
        self.h, self.nshards = h, nshards

    def _e(self):
        return _err(self.lib)

    def apply_events(self, events: Sequence[Event], threads: int = 0) -> None:
        """ncdh_apply_events: a batch of informer events, canonicalised on `threads` threads."""
        if isinstance(events, EventBatch):
            arr, n = events.events, events.n
        else:
            arr, n = (Event * max(len(events), 1))(*events), len(events)
        if self.lib.ncdh_apply_events(self.h, arr, n, threads) != 0:
            raise self._e()

    def diff_digest(self) -> Tuple[int, int]:
        """(layout-independent checksum of the last scan's diff list, objects with at least one entry)."""
        d, n = C.c_uint64(), C.c_uint64()
        if self.lib.ncdh_diff_digest(self.h, C.byref(d), C.byref(n)) != 0:
            raise self._e()
        return d.value, n.value

    def timing(self) -> dict:
        t = ScanTiming()
        if self.lib.ncdh_last_timing(self.h, C.byref(t)) != 0:
            raise self._e()
        return t.as_dict()

    def plan_adoption(self, namespace: str, name: str) -> List[List[str]]:
        """[verb, resource, name] per Update adoptReferences would issue on the controller cluster."""
        n = self.lib.ncdh_plan_adoption(self.h, namespace.encode(), name.encode(), None, 0)
        if n == -STALE:
            raise Stale("stale snapshot")
        if n < 0:
            raise self._e()
        buf = C.create_string_buffer(max(n, 1))
        self.lib.ncdh_plan_adoption(self.h, namespace.encode(), name.encode(), buf, n)
        return [line.split("\t")[1:] for line in buf.raw[:n].decode().splitlines()]

    def close(self):
        if self.h:
            self.lib.ncdh_checker_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def add(self, cluster: int, kind: int, namespace: str, name: str, uid: str, value_wire: bytes,
            owner_refs: Sequence[Tuple[str, str]] = (), secret_names: Sequence[str] = (), configmap_names: Sequence[str] = ()):
        rc = self.lib.ncdh_cluster_add(
            self.h, cluster, kind, namespace.encode(), name.encode(), uid.encode(), value_wire, len(value_wire),
            _strs([k for k, _ in owner_refs]), _strs([u for _, u in owner_refs]), len(owner_refs),
            _strs(list(secret_names)), len(secret_names), _strs(list(configmap_names)), len(configmap_names))
        if rc != 0:
            raise self._e()

    def remove(self, cluster: int, kind: int, namespace: str, name: str) -> bool:
        return self.lib.ncdh_cluster_remove(self.h, cluster, kind, namespace.encode(), name.encode()) == 0

    def scan(self):
        if self.lib.ncdh_scan(self.h) != 0:
            raise self._e()

    def lookup(self, kind: int, namespace: str, name: str, dep: int, shard: int) -> Optional[int]:
        code = C.c_uint32()
        rc = self.lib.ncdh_lookup(self.h, kind, namespace.encode(), name.encode(), dep, shard, C.byref(code))
        if rc < 0:
            raise self._e()
        if rc == STALE:
            raise Stale("stale snapshot")
        return None if rc == NOT_FOUND else code.value

    def _plan(self, fn, namespace: str, name: str):
        n = fn(self.h, namespace.encode(), name.encode(), None, 0)
        if n == -STALE:
            raise Stale("stale snapshot")
        if n < 0:
            raise self._e()
        buf = C.create_string_buffer(max(n, 1))
        fn(self.h, namespace.encode(), name.encode(), buf, n)
        nexus = [[] for _ in range(self.nshards)]
        kube = [[] for _ in range(self.nshards)]
        error = None
        for line in buf.raw[:n].decode().splitlines():
            parts = line.split("\t")
            if parts[0] == "error":
                error = parts[1]
            else:
                (nexus if parts[1] == "nexus" else kube)[int(parts[0])].append([parts[2], parts[3], parts[4]])
        return nexus, kube, error

    def plan_template(self, namespace: str, name: str):
        """(nexus actions per shard, kube actions per shard, error or None): the shard writes
        templateSyncHandler would issue, in order."""
        return self._plan(self.lib.ncdh_plan_template, namespace, name)

    def plan_workgroup(self, namespace: str, name: str):
        return self._plan(self.lib.ncdh_plan_workgroup, namespace, name)

    def sweep_queue(self) -> List[Tuple[str, str]]:
        """(objTypeName, "namespace/name") of every template / workgroup the last scan found work for."""
        n = self.lib.ncdh_sweep_queue(self.h, None, 0)
        if n < 0:
            raise self._e()
        buf = C.create_string_buffer(max(n, 1))
        self.lib.ncdh_sweep_queue(self.h, buf, n)
        return [tuple(line.split("\t")) for line in buf.raw[:n].decode().splitlines()]

    def upload_counts(self, cluster: int) -> Tuple[int, int, int]:
        """(full uploads, structural re-commits, value-only syncs) of one cluster so far."""
        a, b, d = C.c_uint64(), C.c_uint64(), C.c_uint64()
        if self.lib.ncdh_upload_counts(self.h, cluster, C.byref(a), C.byref(b), C.byref(d)) != 0:
            raise self._e()
        return a.value, b.value, d.value

    def counts(self) -> Tuple[int, int]:
        a, b = C.c_uint64(), C.c_uint64()
        if self.lib.ncdh_last_counts(self.h, C.byref(a), C.byref(b)) != 0:
            raise self._e()
        return a.value, b.value
