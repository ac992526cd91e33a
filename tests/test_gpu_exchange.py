"""The peer-memory exchange of diff lists (ncd_xchg_*, ncd_exchange.cu) against the oracle.

One GPU is enough for the kernel path: every "rank" is its own context (own stream, own arenas,
own exchange buffer) on cuda:0 and the ends are connected by device pointer.  The two-process
form (memory handles shipped through a process group) needs two GPUs and is skipped otherwise.
"""
import os
import socket
import sys

import numpy as np
import pytest

import oracle
from nexus_configuration_controller_b200 import arena as A
from nexus_configuration_controller_b200 import multigpu as M
from nexus_configuration_controller_b200._ffi import NcdError
from nexus_configuration_controller_b200.drift import Context
from tests.helpers import random_clusters

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class Rank:
    def __init__(self, ctx, world, rank, seed, cap, n_shards=3):
        self.ctx = ctx
        ctrl, shards = random_clusters(seed, n_templates=12 + 5 * rank, n_shards=n_shards)
        self.want_codes, self.want_ents = oracle.scan(ctrl, shards)
        self.ca = ctx.arena_from_host(ctrl)
        self.sas = [ctx.arena_from_host(h) for h in shards]
        self.x = M.PeerExchange(ctx, world, rank, cap)
        # Ranks share one process and one GPU here: when a scan grows its scratch or allocates
        # digests (cudaFree / cudaMalloc), the driver waits for the whole device, i.e. for another
        # rank's exchange kernel, which waits for this rank.  Separate processes do not have this
        # coupling.  So let every mode allocate what it needs before the first real exchange.
        with M.PeerExchange(ctx, 1, 0, cap) as alone:
            for mode in (A.MODE_DIRECT, A.MODE_HASH64, A.MODE_XXH64):
                alone.scan(self.ca, self.sas, mode)
                alone.scan(self.ca, [], mode)
            alone.fetch()

    def close(self):
        self.x.close()
        for a in self.sas + [self.ca]:
            a.close()


def make_ranks(ctx, world, cap, seeds=None):
    ctxs = [ctx] + [Context(0) for _ in range(world - 1)]
    ranks = [Rank(c, world, r, (seeds or range(20, 20 + world))[r], cap) for r, c in enumerate(ctxs)]
    M.PeerExchange.connect_local([r.x for r in ranks])
    return ctxs, ranks


def close_all(ctxs, ranks):
    for r in ranks:
        r.close()
    for c in ctxs[1:]:
        c.close()


def test_three_ranks_every_rank_holds_every_list(ctx):
    world, cap = 3, 4096
    ctxs, ranks = make_ranks(ctx, world, cap)
    try:
        assert all(len(r.want_ents) > 0 for r in ranks)
        # seven steps: every generation several times, modes alternating; ranks enqueue in a
        # different order each step, so every rank is at times the one that waits
        for step in range(7):
            mode = (A.MODE_DIRECT, A.MODE_HASH64, A.MODE_XXH64)[step % 3]
            order = np.roll(np.arange(world), step)
            for r in order:
                ranks[r].x.scan(ranks[r].ca, ranks[r].sas, mode)
            for r in ranks:
                parts = M.unpack_gathered(r.x.fetch(), world, cap)
                for src, (ns, rows, ents) in enumerate(parts):
                    assert (ns, rows) == (len(ranks[src].sas), ranks[src].want_codes.shape[0]), (step, src)
                    np.testing.assert_array_equal(ents, ranks[src].want_ents)
    finally:
        close_all(ctxs, ranks)


def test_ranks_may_run_one_step_apart(ctx):
    """An exchange waits for the peers' lists of the PREVIOUS step only: a rank finishes its step without the
    slower peer having started the same step, and stalls only when it would get two steps ahead.  The lists of
    a step are complete for whoever fetches them (the fetch waits for that step's flags)."""
    world, cap = 2, 4096
    ctxs, ranks = make_ranks(ctx, world, cap)
    try:
        fast, slow = ranks
        fast.x.scan(fast.ca, fast.sas, A.MODE_DIRECT)  # step 1: nothing to wait for
        fast.ctx.sync()  # returns although the peer has not scanned at all
        slow.x.scan(slow.ca, slow.sas, A.MODE_DIRECT)  # the peer's step 1
        fast.x.scan(fast.ca, fast.sas, A.MODE_DIRECT)  # step 2 waits for the peer's step 1: there already
        fast.ctx.sync()
        slow.x.scan(slow.ca, slow.sas, A.MODE_DIRECT)
        for r in ranks:
            parts = M.unpack_gathered(r.x.fetch(), world, cap)
            for src, (ns, rows, ents) in enumerate(parts):
                np.testing.assert_array_equal(ents, ranks[src].want_ents)
        # ten more steps in lockstep order: every generation is reused several times
        for step in range(10):
            for r in (ranks if step % 2 else ranks[::-1]):
                r.x.scan(r.ca, r.sas, A.MODE_DIRECT)
        for r in ranks:
            parts = M.unpack_gathered(r.x.fetch(), world, cap)
            for src, (ns, rows, ents) in enumerate(parts):
                np.testing.assert_array_equal(ents, ranks[src].want_ents)
    finally:
        close_all(ctxs, ranks)


def test_lists_change_between_steps(ctx):
    """A list that shrinks must not leave stale entries visible: the header carries the count."""
    world, cap = 2, 1024
    ctxs, ranks = make_ranks(ctx, world, cap)
    try:
        for r in ranks:
            r.x.scan(r.ca, r.sas, A.MODE_DIRECT)
        first = [M.unpack_gathered(r.x.fetch(), world, cap) for r in ranks]
        # rank 1 now scans against no shards at all: an empty list
        ranks[0].x.scan(ranks[0].ca, ranks[0].sas, A.MODE_DIRECT)
        ranks[1].x.scan(ranks[1].ca, [], A.MODE_DIRECT)
        for r in ranks:
            parts = M.unpack_gathered(r.x.fetch(), world, cap)
            np.testing.assert_array_equal(parts[0][2], ranks[0].want_ents)
            assert parts[1][0] == 0 and len(parts[1][2]) == 0
        assert len(first[0][1][2]) == len(ranks[1].want_ents) > 0
    finally:
        close_all(ctxs, ranks)


def test_slot_too_small_is_reported_not_overrun(ctx):
    world = 2
    ctxs, ranks = make_ranks(ctx, world, cap=4)
    try:
        for r in ranks:
            assert len(r.want_ents) > 4
            r.x.scan(r.ca, r.sas, A.MODE_DIRECT)
        for r in ranks:
            g = r.x.fetch()
            for src in range(world):
                assert g[src, 0] == len(ranks[src].want_ents) and g[src, 3] == 1
                np.testing.assert_array_equal(g[src, 4 : 4 + 16].reshape(4, 4), ranks[src].want_ents[:4])
            with pytest.raises(ValueError):
                M.unpack_gathered(g, world, 4)
    finally:
        close_all(ctxs, ranks)


def test_world_of_one_is_a_plain_scan(ctx):
    ctxs, ranks = make_ranks(ctx, 1, cap=2048)
    try:
        r = ranks[0]
        for _ in range(3):
            r.x.scan(r.ca, r.sas, A.MODE_DIRECT)
            (ns, rows, ents), = M.unpack_gathered(r.x.fetch(), 1, 2048)
            np.testing.assert_array_equal(ents, r.want_ents)
    finally:
        close_all(ctxs, ranks)


def test_a_peer_that_never_comes_times_out(ctx):
    ctxs, ranks = make_ranks(ctx, 2, cap=256)
    try:
        ranks[0].x.set_timeout(0.2)
        ranks[0].x.scan(ranks[0].ca, ranks[0].sas, A.MODE_DIRECT)
        with pytest.raises(NcdError, match="did not arrive"):
            ranks[0].x.fetch()
    finally:
        close_all(ctxs, ranks)


def test_misuse_is_refused(ctx):
    x = M.PeerExchange(ctx, 2, 0, 16)
    ctrl, shards = random_clusters(3, n_templates=4, n_shards=1)
    ca = ctx.arena_from_host(ctrl)
    try:
        with pytest.raises(NcdError, match="not connected"):
            x.scan(ca, [], A.MODE_DIRECT)
        with pytest.raises(NcdError, match="no exchange"):
            x.fetch()
        with pytest.raises(NcdError):
            M.PeerExchange(ctx, 2, 2, 16)  # rank out of range
        with pytest.raises(NcdError):
            M.PeerExchange(ctx, 0, 0, 16)
        with pytest.raises(NcdError):
            M.PeerExchange(ctx, 1, 0, 0)
        assert len(x.ipc_handle()) == 64
    finally:
        ca.close()
        x.close()


def _two_gpu_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    ok, why = True, ""
    try:
        with Context(rank) as c:
            me = Rank(c, world, rank, 40 + rank, 2048)
            want = [oracle.scan(*random_clusters(40 + r, n_templates=12 + 5 * r, n_shards=3))[1] for r in range(world)]
            me.x.connect_group()
            for step in range(5):
                me.x.scan(me.ca, me.sas, A.MODE_DIRECT)
                parts = M.unpack_gathered(me.x.fetch(), world, 2048)
                for src in range(world):
                    if not np.array_equal(parts[src][2], want[src]):
                        ok, why = False, f"step {step}: list of rank {src} differs on rank {rank}"
            dist.barrier()
            me.close()
    except Exception as e:  # noqa: BLE001 - reported to the parent
        ok, why = False, repr(e)
    q.put((rank, ok, why))
    dist.destroy_process_group()


def test_two_processes_two_gpus():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    procs = [mpc.Process(target=_two_gpu_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert results == [(0, True, ""), (1, True, "")]
