"""The N>1 host logic on CPU: key-hash partition and the single all-gather exchange (gloo, world 2)."""
import os
import socket
import sys

import numpy as np
import pytest

from nexus_configuration_controller_b200 import multigpu as M
from tests import synth_spec

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_is_a_partition_and_matches_the_device_rule():
    keys = [synth_spec.template_key(g).decode() for g in range(2000)]
    for world in (1, 2, 4, 8):
        ranks = [M.owner_rank(0, k, world) for k in keys]
        assert set(ranks) <= set(range(world))
        # same rule as the generator / the kernels (ncd_keyhash % world)
        assert ranks == [synth_spec.keyhash(0, k.encode()) % world if world > 1 else 0 for k in keys]
        if world > 1:
            counts = np.bincount(ranks, minlength=world)
            assert counts.min() > 0.8 * len(keys) / world  # balanced within 20 % at 2000 keys


def test_pack_unpack_round_trip_and_capacity():
    ents = np.array([[1, 0, 2, 1], [5, 2, 0, 6]], dtype=np.uint32)
    buf = M.pack_host(ents, 3, 77, cap_entries=8)
    assert len(buf) == M.buffer_words(8)
    (ns, rows, got), = M.unpack_gathered(buf, 1, 8)
    assert (ns, rows) == (3, 77) and np.array_equal(got, ents)
    with pytest.raises(ValueError):
        M.pack_host(np.zeros((9, 4), dtype=np.uint32), 1, 1, cap_entries=8)


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    cap = 64
    # each rank "scans" its own key range: entries carry the rank in obj so the merge is checkable
    n = 3 + 2 * rank
    ents = np.array([[1000 * rank + i, i % 3, i % 2, 1 + (i % 2)] for i in range(n)], dtype=np.uint32)
    send = torch.from_numpy(M.pack_host(ents, 2, 10 + rank, cap).view(np.int32))
    recv = torch.zeros(world * send.numel(), dtype=torch.int32)
    dist.all_gather_into_tensor(recv, send)  # the one collective of the path
    parts = M.unpack_gathered(recv.numpy().view(np.uint32), world, cap)
    ok = len(parts) == world
    for r, (ns, rows, e) in enumerate(parts):
        ok = ok and ns == 2 and rows == 10 + r and len(e) == 3 + 2 * r and all(int(x) // 1000 == r for x in e[:, 0])
    out.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_two_ranks_exchange_diff_lists_over_gloo():
    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(results) == [(0, True), (1, True)]
