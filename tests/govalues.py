"""A tiny model of Go values for fixtures: enough to say exactly what reflect.DeepEqual sees.

Nil-ness is explicit (a nil map is not an empty map), map entries keep insertion order (so the
canonical encoder's sorting is exercised), integers carry their width.  ``wire()`` serialises a
tree in a plain tag-length-value form that both the oracle (oracle/go_values.cpp) and the
product's host layer (csrc/host) parse; it is a transport for test inputs, not the canonical
encoding.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass
from typing import Any, List, Optional, Sequence, Tuple

T_BOOL, T_INT, T_STRING, T_BYTES, T_NILBYTES, T_SLICE, T_NILSLICE = 1, 2, 3, 4, 5, 6, 7
T_MAP, T_NILMAP, T_PTR, T_NILPTR, T_STRUCT, T_FLOAT, T_IFACE, T_NILIFACE = 8, 9, 10, 11, 12, 13, 14, 15


@dataclass(frozen=True)
class V:
    """One node: kind tag + payload (python value / list of V / list of (V, V) pairs)."""

    t: int
    v: Any = None
    w: int = 0  # integer width for T_INT, type name for T_IFACE lives in `name`
    name: str = ""

    def wire(self) -> bytes:
        t = self.t
        if t == T_BOOL:
            return bytes([t, 1 if self.v else 0])
        if t == T_INT:
            return bytes([t, self.w]) + struct.pack("<q", self.v)
        if t == T_STRING:
            b = self.v.encode() if isinstance(self.v, str) else bytes(self.v)
            return bytes([t]) + struct.pack("<I", len(b)) + b
        if t == T_BYTES:
            return bytes([t]) + struct.pack("<I", len(self.v)) + bytes(self.v)
        if t in (T_NILBYTES, T_NILSLICE, T_NILMAP, T_NILPTR, T_NILIFACE):
            return bytes([t])
        if t in (T_SLICE, T_STRUCT):
            return bytes([t]) + struct.pack("<I", len(self.v)) + b"".join(x.wire() for x in self.v)
        if t == T_MAP:
            return bytes([t]) + struct.pack("<I", len(self.v)) + b"".join(k.wire() + x.wire() for k, x in self.v)
        if t == T_PTR:
            return bytes([t]) + self.v.wire()
        if t == T_FLOAT:
            return bytes([t]) + struct.pack("<d", self.v)
        if t == T_IFACE:
            n = self.name.encode()
            return bytes([t]) + struct.pack("<I", len(n)) + n + self.v.wire()
        raise ValueError(t)

    def to_json(self):
        t = self.t
        if t in (T_BOOL, T_FLOAT):
            return {"t": t, "v": self.v}
        if t == T_INT:
            return {"t": t, "w": self.w, "v": self.v}
        if t == T_STRING:
            return {"t": t, "v": self.v}
        if t == T_BYTES:
            return {"t": t, "v": bytes(self.v).hex()}
        if t in (T_SLICE, T_STRUCT):
            return {"t": t, "v": [x.to_json() for x in self.v]}
        if t == T_MAP:
            return {"t": t, "v": [[k.to_json(), x.to_json()] for k, x in self.v]}
        if t == T_PTR:
            return {"t": t, "v": self.v.to_json()}
        if t == T_IFACE:
            return {"t": t, "name": self.name, "v": self.v.to_json()}
        return {"t": t}

    @staticmethod
    def from_json(j) -> "V":
        t = j["t"]
        if t in (T_BOOL, T_FLOAT, T_STRING):
            return V(t, j["v"])
        if t == T_INT:
            return V(t, j["v"], w=j["w"])
        if t == T_BYTES:
            return V(t, bytes.fromhex(j["v"]))
        if t in (T_SLICE, T_STRUCT):
            return V(t, [V.from_json(x) for x in j["v"]])
        if t == T_MAP:
            return V(t, [(V.from_json(k), V.from_json(x)) for k, x in j["v"]])
        if t == T_PTR:
            return V(t, V.from_json(j["v"]))
        if t == T_IFACE:
            return V(t, V.from_json(j["v"]), name=j["name"])
        return V(t)


def Bool(b: bool) -> V:
    return V(T_BOOL, bool(b))


def Int(v: int, w: int = 64) -> V:
    return V(T_INT, int(v), w=w)


def Int32(v: int) -> V:
    return Int(v, 32)


def Str(s: str) -> V:
    return V(T_STRING, s)


def Bytes(b: Optional[bytes]) -> V:
    return V(T_NILBYTES) if b is None else V(T_BYTES, bytes(b))


def Slice(items: Optional[Sequence[V]]) -> V:
    """None = nil slice, [] = empty non-nil slice (they are NOT DeepEqual)."""
    return V(T_NILSLICE) if items is None else V(T_SLICE, list(items))


def Map(entries: Optional[Sequence[Tuple[V, V]]]) -> V:
    """None = nil map, [] = empty non-nil map (make(map...)); insertion order is kept."""
    return V(T_NILMAP) if entries is None else V(T_MAP, list(entries))


def StrMap(d: Optional[dict]) -> V:
    return Map(None if d is None else [(Str(k), Str(v)) for k, v in d.items()])


def BytesMap(d: Optional[dict]) -> V:
    return Map(None if d is None else [(Str(k), Bytes(v)) for k, v in d.items()])


def Ptr(x: Optional[V]) -> V:
    return V(T_NILPTR) if x is None else V(T_PTR, x)


def Struct(*fields: V) -> V:
    return V(T_STRUCT, list(fields))


def Float(x: float) -> V:
    return V(T_FLOAT, float(x))


def Iface(name: Optional[str], x: Optional[V] = None) -> V:
    return V(T_NILIFACE) if name is None else V(T_IFACE, x, name=name)
