"""Incremental arenas: records rewritten in the pinned host arrays are pulled to the resident
device copy (ncd_arena_sync_records) and the next scan must decide exactly as a fresh full
upload of the same contents would — checked against the oracle on the downloaded arenas."""
import random

import numpy as np
import pytest

import oracle
from nexus_configuration_controller_b200 import arena as A
from nexus_configuration_controller_b200.drift import Arena
from tests.helpers import rand_bytes, random_clusters

pytestmark = pytest.mark.gpu


def resident(ctx, host: A.HostArena, slack: int) -> Arena:
    a = Arena(ctx, host.payload_bytes + slack, host.nobj, host.aux_bytes, host.nowner, host.ndep, device_only=False)
    a.fill_from(host)
    a.commit(host.nobj, host.payload_bytes, host.aux_bytes, host.nowner, host.ndep)
    return a


def rewrite(rng, arr, used: int, i: int, new: bytes) -> int:
    """What the packer does for an UpdateFunc event: in place if the record fits, else appended."""
    old_pad = (int(arr["rec_len"][i]) + 15) // 16 * 16
    new_pad = (len(new) + 15) // 16 * 16
    off = int(arr["rec_off"][i])
    if new_pad > old_pad:
        off, used = used, used + new_pad
    arr["rec_off"][i], arr["rec_len"][i] = off, len(new)
    arr["payload"][off : off + len(new)] = np.frombuffer(new, dtype=np.uint8) if new else []
    arr["payload"][off + len(new) : off + new_pad] = 0
    return used


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_rewritten_records_follow_to_the_device(ctx, seed):
    rng = random.Random(seed)
    ctrl_h, shard_hs = random_clusters(seed + 20, n_templates=14, n_shards=3)
    slack = 1 << 20
    arenas = [resident(ctx, h, slack) for h in [ctrl_h] + shard_hs]
    used = [h.payload_bytes for h in [ctrl_h] + shard_hs]
    try:
        for rnd in range(4):
            for k, a in enumerate(arenas):
                arr = a.host_arrays()
                n = a.counts().nobj
                picks = rng.sample(range(n), max(1, n // 3))
                for i in picks:
                    old = int(arr["rec_len"][i])
                    how = rng.random()
                    if how < 0.35:  # same length, one byte changed
                        new = bytearray(arr["payload"][int(arr["rec_off"][i]) : int(arr["rec_off"][i]) + old].tobytes())
                        if new:
                            new[rng.randrange(len(new))] ^= rng.randint(1, 255)
                        new = bytes(new)
                    elif how < 0.6:  # shorter
                        new = rand_bytes(rng, rng.randint(0, old))
                    else:  # longer: appended behind the payload end
                        new = rand_bytes(rng, old + rng.randint(1, 9000))
                    used[k] = rewrite(rng, arr, used[k], i, new)
                a.sync_records(np.array(picks, dtype=np.uint32), used[k])
            # also make a shard converge on the controller for some objects (drift goes away)
            want_codes, want_ents = oracle.scan(arenas[0].download(), [s.download() for s in arenas[1:]])
            for mode in (A.MODE_DIRECT, A.MODE_HASH64, A.MODE_SHA256):
                d = ctx.scan(arenas[0], arenas[1:], mode)
                np.testing.assert_array_equal(d.codes(), want_codes)
                np.testing.assert_array_equal(d.entries(), want_ents)
                d.close()
            # the device copy is byte-identical to the pinned host arrays where records live
            for a in arenas:
                arr, dev = a.host_arrays(), a.download()
                for i in range(dev.nobj):
                    off, ln = int(arr["rec_off"][i]), int(arr["rec_len"][i])
                    assert dev.record(i) == arr["payload"][off : off + ln].tobytes()
    finally:
        for a in arenas:
            a.close()


def test_sync_refuses_what_it_cannot_place(ctx):
    from nexus_configuration_controller_b200._ffi import NcdError

    host = A.pack([A.CachedObject(A.KIND_WORKGROUP, "default/wg", b"x" * 40)])
    a = resident(ctx, host, 64)
    try:
        with pytest.raises(NcdError):  # beyond capacity
            a.sync_records(np.array([0], dtype=np.uint32), host.payload_bytes + 4096)
        arr = a.host_arrays()
        arr["rec_off"][0] = 8  # not 16-byte aligned
        with pytest.raises(NcdError):
            a.sync_records(np.array([0], dtype=np.uint32), host.payload_bytes)
        with pytest.raises(NcdError):  # the arena needs a commit after a refused sync
            ctx.scan(a, [], A.MODE_DIRECT)
        dev_only = ctx.synth_arena(__import__("tests.test_gpu_parity", fromlist=["small_synth"]).small_synth(groups=4), -1)
        with pytest.raises(NcdError):
            dev_only.sync_records(np.array([0], dtype=np.uint32), 16)
        dev_only.close()
    finally:
        a.close()
