"""Multi-GPU layout of the drift check: partition by object-key hash, one all-gather of diff lists.

One process per GPU.  A template (with its dependents) belongs to rank
``ncd_keyhash(TEMPLATE, "namespace/name") % world``; a workgroup likewise with its own kind.  Every
rank holds the controller copy and all shard copies of its key range, so join and compare need no
cross-GPU traffic.  After the scan each rank packs ``[n, nshards, rows, 0 | n x ncd_entry]`` into a
fixed-size device buffer (``ncd_diff_pack``) and ONE ``all_gather_into_tensor`` hands every rank
every list.  PyTorch is only the plumbing here (process group, NCCL / gloo, streams).
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

from . import _ffi

HEADER_WORDS = 4
ENTRY_WORDS = 4


def owner_rank(kind: int, key: str, world: int) -> int:
    """Rank that owns an object key (host side of the partition; same hash as the device's)."""
    k = key.encode()
    return int(_ffi.load().ncd_keyhash(kind, k, len(k)) % world) if world > 1 else 0


def buffer_words(cap_entries: int) -> int:
    return HEADER_WORDS + ENTRY_WORDS * cap_entries


def pack_host(entries: np.ndarray, nshards: int, rows: int, cap_entries: int) -> np.ndarray:
    """Host-side twin of ncd_diff_pack (used by the CPU tests of the exchange)."""
    n = len(entries)
    if n > cap_entries:
        raise ValueError(f"{n} entries exceed the exchange capacity {cap_entries}")
    buf = np.zeros(buffer_words(cap_entries), dtype=np.uint32)
    buf[:HEADER_WORDS] = [n, nshards, rows & 0xFFFFFFFF, 0]
    buf[HEADER_WORDS : HEADER_WORDS + ENTRY_WORDS * n] = np.asarray(entries, dtype=np.uint32).reshape(-1)
    return buf


def unpack_gathered(gathered: np.ndarray, world: int, cap_entries: int) -> List[Tuple[int, int, np.ndarray]]:
    """Splits the all-gathered buffer into per-rank (nshards, rows, entries[n, 4])."""
    words = buffer_words(cap_entries)
    g = np.asarray(gathered, dtype=np.uint32).reshape(world, words)
    out = []
    for r in range(world):
        n = int(g[r, 0])
        if n > cap_entries:
            raise ValueError(f"rank {r} reports {n} entries, capacity is {cap_entries}")
        out.append((int(g[r, 1]), int(g[r, 2]), g[r, HEADER_WORDS : HEADER_WORDS + ENTRY_WORDS * n].reshape(n, ENTRY_WORDS).copy()))
    return out


def gather_diffs(diff, sendbuf, recvbuf, scan_stream, cap_entries: int, group=None):
    """Enqueues pack + all-gather after the scan that produced `diff`; returns nothing (device work).

    sendbuf / recvbuf: int32 CUDA tensors of buffer_words(cap) and world * buffer_words(cap) words.
    scan_stream: torch.cuda.ExternalStream wrapping the context's stream.  The collective runs on
    torch's current stream, ordered after the scan and before the next one.
    """
    import torch
    import torch.distributed as dist

    diff.pack_into(sendbuf.data_ptr(), cap_entries)
    cur = torch.cuda.current_stream()
    cur.wait_stream(scan_stream)
    dist.all_gather_into_tensor(recvbuf, sendbuf, group=group)
    scan_stream.wait_stream(cur)


def scan_and_gather(ctx, ctrl, shards, mode: int, sendbuf, recvbuf, scan_stream, cap_entries: int, group=None):
    """One multi-GPU step with no host round trip: the scan writes its diff list straight into
    `sendbuf` (ncd_scan_into), the all-gather is queued right behind it.  Everything is enqueued
    when this returns; synchronise before reading `recvbuf`."""
    import torch
    import torch.distributed as dist

    ctx.scan_into(ctrl, shards, mode, sendbuf.data_ptr(), cap_entries)
    if recvbuf is not None:
        cur = torch.cuda.current_stream()
        cur.wait_stream(scan_stream)
        dist.all_gather_into_tensor(recvbuf, sendbuf, group=group)
        scan_stream.wait_stream(cur)


class PeerExchange:
    """One rank's end of the peer-memory exchange of diff lists (``ncd_xchg_*``, ncd_exchange.cu).

    Replaces ``all_gather_into_tensor`` in the step: the compaction kernel writes this rank's list
    into its slot, one more kernel stores the slot into every peer's buffer over NVLink, flags it
    and waits for the peers.  ``torch.distributed`` is used once, to ship the 64-byte memory
    handles (``connect_group``); ranks living in one process connect by pointer (``connect_local``).
    """

    def __init__(self, ctx, world: int, rank: int, cap_entries: int):
        import ctypes as C

        self.ctx, self.world, self.rank, self.cap = ctx, world, rank, cap_entries
        self.lib = ctx.lib
        h = C.c_void_p()
        _ffi.check(self.lib.ncd_xchg_create(ctx.handle, world, rank, cap_entries, C.byref(h)))
        self.handle = h

    # -- set-up
    def ipc_handle(self) -> bytes:
        import ctypes as C

        buf = C.create_string_buffer(64)
        _ffi.check(self.lib.ncd_xchg_handle(self.handle, buf))
        return buf.raw

    def open(self, handles: List[bytes]) -> None:
        assert len(handles) == self.world and all(len(h) == 64 for h in handles)
        _ffi.check(self.lib.ncd_xchg_open(self.handle, b"".join(handles)))

    def connect_group(self, group=None) -> None:
        """Ranks in separate processes: exchange the handles through the process group."""
        import torch.distributed as dist

        if self.world == 1:
            return
        handles: List[bytes] = [b""] * self.world
        dist.all_gather_object(handles, self.ipc_handle(), group=group)
        err = ""
        try:
            self.open(handles)
        except _ffi.NcdError as e:
            err = str(e)
        # doubles as the barrier: nobody pushes before every rank has mapped every buffer, and a
        # rank that could not map its peers fails the set-up on every rank instead of leaving the
        # others waiting
        errs: List[str] = [""] * self.world
        dist.all_gather_object(errs, err, group=group)
        bad = [f"rank {r}: {e}" for r, e in enumerate(errs) if e]
        if bad:
            raise RuntimeError("peer exchange set-up failed: " + "; ".join(bad))

    def buffer(self) -> int:
        import ctypes as C

        p = C.c_void_p()
        _ffi.check(self.lib.ncd_xchg_buffer(self.handle, C.byref(p), None))
        return p.value

    @staticmethod
    def connect_local(ends: "List[PeerExchange]") -> None:
        """Ranks in one process (one context per rank): connect by device pointer."""
        import ctypes as C

        ptrs = (C.c_void_p * len(ends))(*[e.buffer() for e in ends])
        for e in ends:
            if e.world > 1:
                _ffi.check(e.lib.ncd_xchg_connect(e.handle, ptrs))

    def set_timeout(self, seconds: float) -> None:
        _ffi.check(self.lib.ncd_xchg_set_timeout(self.handle, seconds))

    # -- per step
    def scan(self, ctrl, shards, mode: int) -> None:
        """Enqueue scan + exchange on the context's stream (asynchronous)."""
        import ctypes as C

        arr = (C.c_void_p * max(len(shards), 1))(*[s.handle for s in shards])
        _ffi.check(self.lib.ncd_scan_exchange(self.ctx.handle, ctrl.handle, arr, len(shards), mode, self.handle))

    def fetch(self) -> np.ndarray:
        """Waits for the latest exchange and returns uint32[world, buffer_words(cap)] (see unpack_gathered)."""
        out = np.zeros(self.world * buffer_words(self.cap), dtype=np.uint32)
        _ffi.check(self.lib.ncd_xchg_fetch(self.handle, out.ctypes.data, out.nbytes))
        return out.reshape(self.world, buffer_words(self.cap))

    def close(self) -> None:
        if self.handle:
            self.lib.ncd_xchg_destroy(self.handle)
            self.handle = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
