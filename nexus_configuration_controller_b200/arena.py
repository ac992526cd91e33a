"""Host-side arena images: the structure-of-arrays layout of include/ncdrift.h as numpy arrays.

An arena is one cluster's informer cache in canonical form (what the listers of
controller.go:268-278 / shards.Shard hold, reduced to what the drift check reads).  This module
only lays bytes out; it computes nothing about drift.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Sequence, Tuple

import numpy as np

from . import _ffi

KIND_TEMPLATE, KIND_WORKGROUP, KIND_SECRET, KIND_CONFIGMAP = 0, 1, 2, 3
KIND_NAMES = {0: "template", 1: "workgroup", 2: "secret", 3: "configmap"}

SKIP, CREATE, UPDATE_DATA, UPDATE_OWNER, ERR_ROGUE, ERR_MISSING_REF = 0, 1, 2, 4, 8, 16
CODE_NAMES = {
    0: "SKIP",
    1: "CREATE",
    2: "UPDATE_DATA",
    4: "UPDATE_OWNER",
    6: "UPDATE_DATA|UPDATE_OWNER",
    8: "ERR_ROGUE",
    16: "ERR_MISSING_REF",
}

MODE_DIRECT, MODE_HASH64, MODE_SHA256, MODE_XXH64, MODE_CONFIRM, MODE_REHASH, MODE_CACHED = 0, 1, 2, 3, 0x100, 0x200, 0x400


@dataclass
class CachedObject:
    """One object of an informer cache, already canonicalised.

    key      "namespace/name" (cache.ObjectName.String())
    canon    canonical encoding of the compared value (.Spec or .Data)
    owners   metadata.ownerReferences as (Kind == "NexusAlgorithmTemplate", UID)
    deps     templates only: (kind, "namespace/name") of GetSecretNames() then GetConfigMapNames()
    """

    kind: int
    key: str
    canon: bytes
    uid: str = ""
    owners: List[Tuple[bool, str]] = field(default_factory=list)
    deps: List[Tuple[int, str]] = field(default_factory=list)


class HostArena:
    """numpy image of one arena + a ctypes ncd_arena_view over it."""

    FIELDS = [
        ("payload", np.uint8),
        ("rec_off", np.uint64),
        ("rec_len", np.uint32),
        ("kind", np.uint8),
        ("key_off", np.uint32),
        ("key_len", np.uint32),
        ("uid_off", np.uint32),
        ("uid_len", np.uint32),
        ("own_start", np.uint32),
        ("own_is_template", np.uint8),
        ("own_uid_off", np.uint32),
        ("own_uid_len", np.uint32),
        ("dep_start", np.uint32),
        ("dep_kind", np.uint8),
        ("dep_key_off", np.uint32),
        ("dep_key_len", np.uint32),
        ("aux", np.uint8),
    ]

    def __init__(self, nobj: int, payload_bytes: int, aux_bytes: int, nowner: int, ndep: int):
        self.nobj, self.payload_bytes, self.aux_bytes = nobj, payload_bytes, aux_bytes
        self.nowner, self.ndep = nowner, ndep
        sizes = {
            "payload": max(payload_bytes, 16),
            "aux": max(aux_bytes, 1),
            "own_start": nobj + 1,
            "dep_start": nobj + 1,
            "own_is_template": max(nowner, 1),
            "own_uid_off": max(nowner, 1),
            "own_uid_len": max(nowner, 1),
            "dep_kind": max(ndep, 1),
            "dep_key_off": max(ndep, 1),
            "dep_key_len": max(ndep, 1),
        }
        for name, dt in self.FIELDS:
            setattr(self, name, np.zeros(sizes.get(name, max(nobj, 1)), dtype=dt))

    def view(self) -> _ffi.ArenaView:
        v = _ffi.ArenaView()
        v.nobj, v.payload_bytes, v.aux_bytes = self.nobj, self.payload_bytes, self.aux_bytes
        v.nowner, v.ndep = self.nowner, self.ndep
        for name, _ in self.FIELDS:
            arr = getattr(self, name)
            ctype = dict(_ffi.ArenaView._fields_)[name]
            setattr(v, name, arr.ctypes.data_as(ctype))
        return v

    def record(self, i: int) -> bytes:
        off, ln = int(self.rec_off[i]), int(self.rec_len[i])
        return self.payload[off : off + ln].tobytes()

    def key(self, i: int) -> str:
        off, ln = int(self.key_off[i]), int(self.key_len[i])
        return self.aux[off : off + ln].tobytes().decode()


def pack(objects: Sequence[CachedObject]) -> HostArena:
    """Lays a list of cached objects out as an arena (records 16-byte aligned, zero padded)."""
    n = len(objects)
    pad = lambda x: (x + 15) // 16 * 16  # noqa: E731
    payload_bytes = sum(pad(len(o.canon)) for o in objects)
    aux = bytearray()
    nowner = sum(len(o.owners) for o in objects)
    ndep = sum(len(o.deps) for o in objects)

    def put(b: bytes) -> Tuple[int, int]:
        off = len(aux)
        aux.extend(b)
        return off, len(b)

    staged = []
    for o in objects:
        k = put(o.key.encode())
        u = put(o.uid.encode())
        ow = [(t, put(uid.encode())) for t, uid in o.owners]
        dp = [(kd, put(key.encode())) for kd, key in o.deps]
        staged.append((k, u, ow, dp))
    a = HostArena(n, payload_bytes, len(aux), nowner, ndep)
    a.aux[: len(aux)] = np.frombuffer(bytes(aux), dtype=np.uint8) if aux else []
    off = 0
    oi = di = 0
    for i, (o, (k, u, ow, dp)) in enumerate(zip(objects, staged)):
        a.rec_off[i], a.rec_len[i], a.kind[i] = off, len(o.canon), o.kind
        if o.canon:
            a.payload[off : off + len(o.canon)] = np.frombuffer(o.canon, dtype=np.uint8)
        off += pad(len(o.canon))
        a.key_off[i], a.key_len[i] = k
        a.uid_off[i], a.uid_len[i] = u
        a.own_start[i] = oi
        for is_t, (uo, ul) in ow:
            a.own_is_template[oi], a.own_uid_off[oi], a.own_uid_len[oi] = int(is_t), uo, ul
            oi += 1
        a.dep_start[i] = di
        for kd, (ko, kl) in dp:
            a.dep_kind[di], a.dep_key_off[di], a.dep_key_len[di] = kd, ko, kl
            di += 1
    a.own_start[n], a.dep_start[n] = oi, di
    return a


def rows_of(a: HostArena) -> List[Tuple[int, int]]:
    """(obj, dep) of every row a controller arena produces, in order (dep 0 = the object itself)."""
    rows = []
    for i in range(a.nobj):
        if a.kind[i] == KIND_TEMPLATE:
            rows.append((i, 0))
            rows.extend((i, j + 1) for j in range(int(a.dep_start[i + 1] - a.dep_start[i])))
        elif a.kind[i] == KIND_WORKGROUP:
            rows.append((i, 0))
    return rows
