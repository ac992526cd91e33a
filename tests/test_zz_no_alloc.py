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

import oracle
from nexus_configuration_controller_b200 import _ffi
from nexus_configuration_controller_b200 import arena as A
from tests.helpers import rand_bytes

pytestmark = pytest.mark.gpu


def _records(seed, lens):
    rng = random.Random(seed)
    return A.pack([A.CachedObject(A.KIND_SECRET, f"default/s{i}", rand_bytes(rng, n)) for i, n in enumerate(lens)])


def _static_digests(lib, torch, host, stream=None, check_no_alloc=False):
    dev = torch.device("cuda", 0)
    payload = torch.from_numpy(host.payload[: max(host.payload_bytes, 16)].copy()).to(dev)
    rec_off = torch.from_numpy(host.rec_off[: host.nobj].astype(np.int64)).to(dev)
    rec_len = torch.from_numpy(host.rec_len[: host.nobj].astype(np.int32)).to(dev)
    out = torch.full((host.nobj,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    rc = lib.ncd_hash64_static(payload.data_ptr(), rec_off.data_ptr(), rec_len.data_ptr(), host.nobj, out.data_ptr(),
                               C.c_void_p(stream.cuda_stream) if stream is not None else None)
    assert rc == 0, lib.ncd_last_error()
    torch.cuda.synchronize()
    if check_no_alloc:
        # the point of the entry point: a call on a device it has been used on before takes no
        # device memory (nothing was allocated, so the driver's free figure does not move)
        assert torch.cuda.mem_get_info(0)[0] == free_before
    return out.cpu().numpy().view(np.uint64)


def test_static_hash_matches_arena_hash_on_two_calls_in_a_row(ctx):
    import torch

    lib = _ffi.load()
    # first call: many records, several of them spanning tiles; second call: fewer and different
    # records, so anything the first call left in the shared array would show up in the digests
    first = _records(1, [0, 1, 15, 16, 17, 4095, 4096, 4097, 8192, 12289, 70000, 1 << 20] + [random.Random(5).randint(1, 9000) for _ in range(300)])
    second = _records(2, [70001, 3, 4096, 33, 250000] + [random.Random(6).randint(1, 20000) for _ in range(40)])
    for k, host in enumerate((first, second, first)):
        a = ctx.arena_from_host(host)
        want, _ = ctx.hash_arena(a, A.MODE_HASH64)
        a.close()
        got = _static_digests(lib, torch, host, check_no_alloc=k > 0)
        np.testing.assert_array_equal(got, want)
        # and against the CPU statement of NCD-H64 directly, not only through the other GPU path
        np.testing.assert_array_equal(got, oracle.h64_arena(host))


def test_static_hash_on_a_side_stream_with_empty_records_at_the_end():
    """A caller's own (non-default) stream, and records of length zero in the middle and at the end of
    the list: a zero-length record owns no tile, so the tile -> record search has to step over it."""
    import torch

    lib = _ffi.load()
    host = _records(3, [4097, 0, 0, 17, 8192, 0, 5000, 0, 0])
    side = torch.cuda.Stream(device=0)
    want = oracle.h64_arena(host)
    for _ in range(2):
        np.testing.assert_array_equal(_static_digests(lib, torch, host, stream=side), want)
    only_empty = _records(4, [0, 0, 0])
    np.testing.assert_array_equal(_static_digests(lib, torch, only_empty, stream=side), oracle.h64_arena(only_empty))


def test_static_hash_at_its_record_limit():
    """Exactly NCD_STATIC_MAX_RECORDS records in one call: the largest list the fixed scratch holds."""
    import torch

    lib = _ffi.load()
    rng = random.Random(8)
    host = _records(9, [rng.choice([0, 1, 16, 40, 100, 4097]) for _ in range(65536)])
    got = _static_digests(lib, torch, host, check_no_alloc=True)
    np.testing.assert_array_equal(got, oracle.h64_arena(host))


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
