"""Error behaviour of the C ABI: every misuse is a status code + a message, never a crash or a
silent downgrade (the Go caller turns them into `error` and the work queue requeues)."""
import ctypes as C

import numpy as np
import pytest

from nexus_configuration_controller_b200 import _ffi
from nexus_configuration_controller_b200 import arena as A
from nexus_configuration_controller_b200._ffi import NcdError
from nexus_configuration_controller_b200.drift import Arena, Context

pytestmark = pytest.mark.gpu


def one(key="default/wg", data=b"abc"):
    return A.pack([A.CachedObject(A.KIND_WORKGROUP, key, data)])


def test_bad_devices_and_modes(ctx):
    lib = _ffi.load()
    h = C.c_void_p()
    assert lib.ncd_init(9999, C.byref(h)) == -1 and b"out of range" in lib.ncd_last_error()
    a = ctx.arena_from_host(one())
    with pytest.raises(NcdError) as e:
        ctx.scan(a, [a], 7)
    assert e.value.status == -1 and "mode" in str(e.value)
    a.close()


def test_uncommitted_and_foreign_arenas_are_refused(ctx):
    host = one()
    raw = Arena(ctx, host.payload_bytes, host.nobj, host.aux_bytes, 0, 0, device_only=False)  # never committed
    with pytest.raises(NcdError) as e:
        ctx.scan(raw, [], A.MODE_DIRECT)
    assert e.value.status == -5
    good = ctx.arena_from_host(host)
    with pytest.raises(NcdError) as e:
        ctx.scan(good, [raw], A.MODE_DIRECT)
    assert e.value.status == -5 and "shard arena 0" in str(e.value)
    with Context(0) as other:
        with pytest.raises(NcdError):
            other.scan(good, [], A.MODE_DIRECT)  # an arena belongs to the context it was made in
    raw.close()
    good.close()


def test_commit_validates_what_the_packer_wrote(ctx):
    host = one(data=b"x" * 100)
    a = Arena(ctx, host.payload_bytes, host.nobj, host.aux_bytes, 0, 0, device_only=False)
    a.fill_from(host)
    arr = a.host_arrays()
    arr["rec_len"][0] = 10_000  # record runs past the payload
    with pytest.raises(NcdError) as e:
        a.commit(host.nobj, host.payload_bytes, host.aux_bytes, 0, 0)
    assert e.value.status == -1 and "outside the arena" in str(e.value)
    arr["rec_len"][0] = 100
    arr["kind"][0] = 9
    with pytest.raises(NcdError):
        a.commit(host.nobj, host.payload_bytes, host.aux_bytes, 0, 0)
    arr["kind"][0] = A.KIND_WORKGROUP
    with pytest.raises(NcdError) as e:  # more than the arena was created for
        a.commit(host.nobj + 1, host.payload_bytes, host.aux_bytes, 0, 0)
    assert e.value.status == -4
    a.commit(host.nobj, host.payload_bytes, host.aux_bytes, 0, 0)  # and the right contents go through
    d = ctx.scan(a, [a], A.MODE_DIRECT)
    assert d.codes().tolist() == [[A.SKIP]]
    d.close()
    a.close()


def test_one_pending_scan_at_a_time_and_small_output_buffers(ctx):
    a = ctx.arena_from_host(one())
    ctx.scan_launch(a, [a], A.MODE_DIRECT)
    with pytest.raises(NcdError) as e:
        ctx.scan_launch(a, [a], A.MODE_DIRECT)
    assert e.value.status == -5
    d = ctx.scan_finish()
    with pytest.raises(NcdError):
        ctx.scan_finish()
    lib = _ffi.load()
    assert lib.ncd_diff_codes(d.handle, None, 0) == -4  # one code needed, room for none
    out = np.zeros(8, dtype=np.uint8)
    assert lib.ncd_hash_arena(ctx.handle, a.handle, A.MODE_DIRECT, out.ctypes.data_as(C.c_void_p), 8, None) == -1
    assert lib.ncd_hash_arena(ctx.handle, a.handle, A.MODE_SHA256, out.ctypes.data_as(C.c_void_p), 8, None) == -4
    d.close()
    a.close()


def test_duplicate_keys_keep_the_first_object(ctx):
    """Two objects with one key cannot exist in an informer store; if a packer writes them anyway
    the first wins (as in the oracle's store), instead of an arbitrary one."""
    import oracle

    ctrl = A.pack([A.CachedObject(A.KIND_WORKGROUP, "default/wg", b"aaaa")])
    shard = A.pack([A.CachedObject(A.KIND_WORKGROUP, "default/wg", b"aaaa"), A.CachedObject(A.KIND_WORKGROUP, "default/wg", b"bbbb")])
    shard2 = A.pack([A.CachedObject(A.KIND_WORKGROUP, "default/wg", b"bbbb"), A.CachedObject(A.KIND_WORKGROUP, "default/wg", b"aaaa")])
    want, _ = oracle.scan(ctrl, [shard, shard2])
    ca, s1, s2 = ctx.arena_from_host(ctrl), ctx.arena_from_host(shard), ctx.arena_from_host(shard2)
    d = ctx.scan(ca, [s1, s2], A.MODE_DIRECT)
    np.testing.assert_array_equal(d.codes(), want)
    assert d.codes().tolist() == [[A.SKIP, A.UPDATE_DATA]]
    d.close()
    for x in (ca, s1, s2):
        x.close()
