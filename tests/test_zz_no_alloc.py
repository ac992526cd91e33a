"""ncd_hash64_static — the record hash for callers that cannot allocate.

Some users of libncdrift.so run where GPU memory may not be requested after start-up (a GPU
shared under a fixed reservation, a process whose allocator owns the whole device, a handler
that cannot take cudaMalloc's device-wide stall).  They pass their own device buffers and the
library's working memory is a fixed array inside its image.  This test plays such a caller:
the buffers are torch tensors, the call goes through the C ABI, and the digests must be
bit-identical (integer work: tolerance zero) to ncd_hash_arena's on the same records — on two
calls in a row, because the second call finds the first call's leftovers in the shared array.

Kept in a file that sorts last so that a run stopped at the first failure has already run every
other test; it passes alone in a fresh process as well as after the rest of the suite.
"""
import ctypes as C
import random

import numpy as np
import pytest

from nexus_configuration_controller_b200 import _ffi
from nexus_configuration_controller_b200 import arena as A
from tests.helpers import rand_bytes

pytestmark = pytest.mark.gpu


def _records(seed, lens):
    rng = random.Random(seed)
    return A.pack([A.CachedObject(A.KIND_SECRET, f"default/s{i}", rand_bytes(rng, n)) for i, n in enumerate(lens)])


def _static_digests(lib, torch, host):
    dev = torch.device("cuda", 0)
    payload = torch.from_numpy(host.payload[: max(host.payload_bytes, 16)].copy()).to(dev)
    rec_off = torch.from_numpy(host.rec_off[: host.nobj].astype(np.int64)).to(dev)
    rec_len = torch.from_numpy(host.rec_len[: host.nobj].astype(np.int32)).to(dev)
    out = torch.full((host.nobj,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    rc = lib.ncd_hash64_static(payload.data_ptr(), rec_off.data_ptr(), rec_len.data_ptr(), host.nobj, out.data_ptr(), None)
    assert rc == 0, lib.ncd_last_error()
    torch.cuda.synchronize()
    return out.cpu().numpy().view(np.uint64)


def test_static_hash_matches_arena_hash_on_two_calls_in_a_row(ctx):
    import torch

    lib = _ffi.load()
    # first call: many records, several of them spanning tiles; second call: fewer and different
    # records, so anything the first call left in the shared array would show up in the digests
    first = _records(1, [0, 1, 15, 16, 17, 4095, 4096, 4097, 8192, 12289, 70000, 1 << 20] + [random.Random(5).randint(1, 9000) for _ in range(300)])
    second = _records(2, [70001, 3, 4096, 33, 250000] + [random.Random(6).randint(1, 20000) for _ in range(40)])
    for host in (first, second, first):
        a = ctx.arena_from_host(host)
        want, _ = ctx.hash_arena(a, A.MODE_HASH64)
        a.close()
        got = _static_digests(lib, torch, host)
        np.testing.assert_array_equal(got, want)


def test_static_hash_refuses_more_records_than_its_scratch_holds():
    import torch

    lib = _ffi.load()
    dummy = torch.zeros(16, dtype=torch.int64, device="cuda:0")
    too_many = 65536 + 1
    rc = lib.ncd_hash64_static(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), too_many, dummy.data_ptr(), None)
    assert rc == -4  # NCD_E_CAPACITY
    msg = lib.ncd_last_error().decode()
    assert "65536" in msg and "65537" in msg
    # nothing was launched, nothing was written
    torch.cuda.synchronize()
    assert int(dummy.abs().sum()) == 0
