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
    launches = {}
    try:
        for rnd in range(4):
            batch = []
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
                if rnd % 2 == 0:
                    a.sync_records(np.array(picks, dtype=np.uint32), used[k])
                else:
                    batch.append(np.array(picks, dtype=np.uint32))
            if batch:  # odd rounds: every arena in one call (ncd_arenas_sync_records)
                ctx.sync_records_many(arenas, batch, used)
            # also make a shard converge on the controller for some objects (drift goes away)
            want_codes, want_ents = oracle.scan(arenas[0].download(), [s.download() for s in arenas[1:]])
            for mode in (A.MODE_DIRECT, A.MODE_HASH64, A.MODE_SHA256, A.MODE_XXH64):
                d = ctx.scan(arenas[0], arenas[1:], mode)
                np.testing.assert_array_equal(d.codes(), want_codes)
                np.testing.assert_array_equal(d.entries(), want_ents)
                d.close()
                # digests survive a value sync (only the rewritten records are hashed again, inside
                # the sync), so from the second round on a hash-mode scan launches no hash kernel
                launches.setdefault(mode, []).append(ctx.last_stats().launches)
            # ... and the digests the arenas hold are those of the new contents
            for a in arenas:
                dev = a.download()
                h64, ms = ctx.hash_arena(a, A.MODE_HASH64 | A.MODE_CACHED)
                np.testing.assert_array_equal(h64, oracle.h64_arena(dev))
                sha, _ = ctx.hash_arena(a, A.MODE_SHA256 | A.MODE_CACHED)
                np.testing.assert_array_equal(sha, oracle.sha256_arena(dev))
                xxh, _ = ctx.hash_arena(a, A.MODE_XXH64 | A.MODE_CACHED)
                assert [int(v) for v in xxh] == [oracle.xxh64(dev.record(i)) for i in range(dev.nobj)]
            # the device copy is byte-identical to the pinned host arrays where records live
            for a in arenas:
                arr, dev = a.host_arrays(), a.download()
                for i in range(dev.nobj):
                    off, ln = int(arr["rec_off"][i]), int(arr["rec_len"][i])
                    assert dev.record(i) == arr["payload"][off : off + ln].tobytes()
        for mode in (A.MODE_HASH64, A.MODE_SHA256, A.MODE_XXH64):
            first, later = launches[mode][0], launches[mode][1:]
            assert all(n < first for n in later) and len(set(later)) == 1, (mode, launches[mode])
            assert later[0] <= launches[A.MODE_DIRECT][-1]  # no hash kernel, and no compare kernel either
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


def test_batched_sync_finishes_the_good_arenas_and_reports_the_bad_one(ctx):
    """ncd_arenas_sync_records with one arena naming an object it does not have, while both hold
    digests: the good arena is synced (and its digests follow), the bad one is refused and needs a
    commit, nothing reads outside an arena."""
    from nexus_configuration_controller_b200._ffi import NcdError

    ctrl_h, shard_hs = random_clusters(9, n_templates=8, n_shards=2)
    good, bad = resident(ctx, shard_hs[0], 1 << 16), resident(ctx, shard_hs[1], 1 << 16)
    try:
        for a in (good, bad):
            ctx.hash_arena(a, A.MODE_HASH64)
            ctx.hash_arena(a, A.MODE_XXH64)
            ctx.hash_arena(a, A.MODE_SHA256)
        arr = good.host_arrays()
        i = next(k for k in range(shard_hs[0].nobj) if shard_hs[0].rec_len[k] > 0)
        arr["payload"][int(arr["rec_off"][i])] ^= 0x11
        with pytest.raises(NcdError, match="refused"):
            ctx.sync_records_many([good, bad], [np.array([i], dtype=np.uint32), np.array([0, 10**6], dtype=np.uint32)],
                                  [shard_hs[0].payload_bytes, shard_hs[1].payload_bytes])
        dev = good.download()
        assert dev.record(i) != shard_hs[0].record(i)
        h64, _ = ctx.hash_arena(good, A.MODE_HASH64 | A.MODE_CACHED)
        np.testing.assert_array_equal(h64, oracle.h64_arena(dev))
        with pytest.raises(NcdError):
            ctx.scan(bad, [], A.MODE_DIRECT)
        with pytest.raises(NcdError):  # the same arena twice in one batch
            ctx.sync_records_many([good, good], [np.array([i], dtype=np.uint32)] * 2, [shard_hs[0].payload_bytes] * 2)
        ctx.sync_records_many([], [], [])  # nothing to do is not an error
    finally:
        good.close()
        bad.close()


def test_structural_change_recommits_without_the_old_payload(ctx):
    """Objects deleted, added and re-owned: the small arrays are rewritten and re-uploaded, records
    that survive stay where they are on the device, only appended records cross PCIe."""
    rng = random.Random(77)
    ctrl_h, shard_hs = random_clusters(31, n_templates=12, n_shards=2)
    arenas = [ctx.arena_from_host(ctrl_h)]
    # shard 0 is resident with room to grow
    sh = shard_hs[0]
    a = Arena(ctx, sh.payload_bytes + (1 << 16), sh.nobj + 8, sh.aux_bytes + 4096, sh.nowner + 16, sh.ndep + 16, device_only=False)
    a.fill_from(sh)
    a.commit(sh.nobj, sh.payload_bytes, sh.aux_bytes, sh.nowner, sh.ndep)
    other = ctx.arena_from_host(shard_hs[1])
    try:
        # scan twice before the change: the second scan answers the join from the cache the first
        # one filled, and the re-commit below must invalidate that cache
        for _ in range(2):
            d = ctx.scan(arenas[0], [a, other], A.MODE_DIRECT)
            np.testing.assert_array_equal(d.codes(), oracle.scan(ctrl_h, shard_hs)[0])
            d.close()
        # rebuild shard 0's object list on the host: drop two objects, add one the controller has
        # but the shard lacked, give one dependent another owner
        objs = []
        for i in range(sh.nobj):
            owners = [(bool(sh.own_is_template[o]), sh.aux[int(sh.own_uid_off[o]) : int(sh.own_uid_off[o]) + int(sh.own_uid_len[o])].tobytes().decode())
                      for o in range(int(sh.own_start[i]), int(sh.own_start[i + 1]))]
            uid = sh.aux[int(sh.uid_off[i]) : int(sh.uid_off[i]) + int(sh.uid_len[i])].tobytes().decode()
            objs.append((A.CachedObject(int(sh.kind[i]), sh.key(i), sh.record(i), uid=uid, owners=owners), int(sh.rec_off[i])))
        del objs[3], objs[0]
        for k, (o, off) in enumerate(objs):
            if o.kind in (A.KIND_SECRET, A.KIND_CONFIGMAP) and o.owners:
                o.owners = [(True, "a-new-owner")]
                break
        have = {(o.kind, o.key) for o, _ in objs}
        added = None
        for i in range(ctrl_h.nobj):
            if (int(ctrl_h.kind[i]), ctrl_h.key(i)) not in have and int(ctrl_h.kind[i]) in (A.KIND_SECRET, A.KIND_CONFIGMAP):
                added = A.CachedObject(int(ctrl_h.kind[i]), ctrl_h.key(i), ctrl_h.record(i), uid="fresh", owners=[(True, "x")])
                break
        assert added is not None
        # write the new image into the pinned arrays: metadata from scratch, surviving records keep
        # their offsets, the added record is appended behind the old payload end
        arr = a.host_arrays()
        dirty_from = sh.payload_bytes
        used = dirty_from
        aux = bytearray()

        def put(b):
            off = len(aux)
            aux.extend(b)
            return off, len(b)

        oi = 0
        final = objs + [(added, None)]
        for i, (o, off) in enumerate(final):
            if off is None:
                off = used
                pad = (len(o.canon) + 15) // 16 * 16
                arr["payload"][off : off + len(o.canon)] = np.frombuffer(o.canon, dtype=np.uint8)
                arr["payload"][off + len(o.canon) : off + pad] = 0
                used += pad
            arr["rec_off"][i], arr["rec_len"][i], arr["kind"][i] = off, len(o.canon), o.kind
            arr["key_off"][i], arr["key_len"][i] = put(o.key.encode())
            arr["uid_off"][i], arr["uid_len"][i] = put(o.uid.encode())
            arr["own_start"][i] = oi
            for is_t, uid in o.owners:
                arr["own_is_template"][oi] = int(is_t)
                arr["own_uid_off"][oi], arr["own_uid_len"][oi] = put(uid.encode())
                oi += 1
            arr["dep_start"][i] = 0
        n = len(final)
        arr["own_start"][n], arr["dep_start"][n] = oi, 0
        arr["aux"][: len(aux)] = np.frombuffer(bytes(aux), dtype=np.uint8)
        a.recommit(n, used, len(aux), oi, 0, dirty_from)
        want_codes, want_ents = oracle.scan(arenas[0].download(), [a.download(), other.download()])
        ref = A.pack([o for o, _ in final])  # the same contents, freshly packed
        ref_codes, _ = oracle.scan(ctrl_h, [ref, shard_hs[1]])
        np.testing.assert_array_equal(want_codes, ref_codes)
        for mode in (A.MODE_DIRECT, A.MODE_HASH64):
            d = ctx.scan(arenas[0], [a, other], mode)
            np.testing.assert_array_equal(d.codes(), ref_codes)
            np.testing.assert_array_equal(d.entries(), want_ents)
            d.close()
    finally:
        for x in arenas + [a, other]:
            x.close()


def test_digests_follow_an_arena_that_grows_between_commits(ctx):
    """An arena made with spare object capacity is hashed while it holds few objects, then committed
    again with more (an informer AddFunc that still fits: ncd_arena_commit / ncd_arena_recommit inside
    nobj_cap).  The digest buffers the arena keeps must be large enough for the later commit: every
    digest kind is checked against the oracle after the growth, and the digests of a neighbouring
    arena allocated right after the first hash stay intact."""
    rng = random.Random(404)
    objs = [A.CachedObject(A.KIND_SECRET, f"default/s{i}", rand_bytes(rng, rng.randint(1, 6000))) for i in range(600)]
    small, big = A.pack(objs[:40]), A.pack(objs)
    a = Arena(ctx, big.payload_bytes + 4096, big.nobj + 8, big.aux_bytes + 64, big.nowner + 1, big.ndep + 1, device_only=False)
    try:
        a.fill_from(small)
        a.commit(small.nobj, small.payload_bytes, small.aux_bytes, small.nowner, small.ndep)
        for mode in (A.MODE_HASH64, A.MODE_SHA256, A.MODE_XXH64):
            ctx.hash_arena(a, mode)
        # a neighbour whose own digests would sit right behind a too-small buffer
        nb = ctx.arena_from_host(small)
        nb_want = {m: ctx.hash_arena(nb, m)[0].copy() for m in (A.MODE_HASH64, A.MODE_SHA256, A.MODE_XXH64)}
        a.fill_from(big)
        a.commit(big.nobj, big.payload_bytes, big.aux_bytes, big.nowner, big.ndep)
        dev = a.download()
        h64, _ = ctx.hash_arena(a, A.MODE_HASH64)
        np.testing.assert_array_equal(h64, oracle.h64_arena(dev))
        sha, _ = ctx.hash_arena(a, A.MODE_SHA256)
        np.testing.assert_array_equal(sha, oracle.sha256_arena(dev))
        xxh, _ = ctx.hash_arena(a, A.MODE_XXH64)
        assert [int(v) for v in xxh] == [oracle.xxh64(dev.record(i)) for i in range(dev.nobj)]
        for m, want in nb_want.items():
            np.testing.assert_array_equal(ctx.hash_arena(nb, m | A.MODE_CACHED)[0], want)
        nb.close()
    finally:
        a.close()


def test_digests_survive_a_recommit(ctx):
    """Objects deleted, renumbered and added by ncd_arena_recommit: the digests of the records that stay where
    they are move with them, only appended records are hashed again, and a hash-mode scan right after the
    re-commit launches no hash kernel (same launch count as the steady state before it)."""
    rng = random.Random(99)
    objs = [A.CachedObject(A.KIND_WORKGROUP, f"default/wg-{i}", rand_bytes(rng, rng.choice([0, 1, 16, 100, 4097, 9000])), uid="u")
            for i in range(300)]
    host = A.pack(objs)
    ctrl = ctx.arena_from_host(host)
    a = Arena(ctx, host.payload_bytes + (1 << 18), host.nobj + 64, host.aux_bytes + 4096, host.nowner + 1, host.ndep + 1, device_only=False)
    try:
        a.fill_from(host)
        a.commit(host.nobj, host.payload_bytes, host.aux_bytes, host.nowner, host.ndep)
        modes = (A.MODE_HASH64, A.MODE_SHA256, A.MODE_XXH64)
        steady = {}
        for m in modes:
            for _ in range(2):
                d = ctx.scan(ctrl, [a], m)
                assert len(d.entries()) == 0
                d.close()
            steady[m] = ctx.last_stats().launches
        # new image: drop some objects, reverse the order of the survivors, append new ones behind the old payload
        keep = [i for i in range(host.nobj) if i % 7 != 3][::-1]
        added = [A.CachedObject(A.KIND_WORKGROUP, f"default/new-{k}", rand_bytes(rng, 50 + 1000 * k), uid="u") for k in range(5)]
        arr = a.host_arrays()
        used, aux = host.payload_bytes, bytearray()

        def put(b):
            off = len(aux)
            aux.extend(b)
            return off, len(b)

        final = []
        for n, i in enumerate(keep):
            arr["rec_off"][n], arr["rec_len"][n], arr["kind"][n] = host.rec_off[i], host.rec_len[i], host.kind[i]
            arr["key_off"][n], arr["key_len"][n] = put(host.key(i).encode())
            arr["uid_off"][n], arr["uid_len"][n] = put(b"u")
            final.append(host.record(i))
        for o in added:
            n = len(final)
            pad = (len(o.canon) + 15) // 16 * 16
            arr["payload"][used : used + len(o.canon)] = np.frombuffer(o.canon, dtype=np.uint8)
            arr["payload"][used + len(o.canon) : used + pad] = 0
            arr["rec_off"][n], arr["rec_len"][n], arr["kind"][n] = used, len(o.canon), o.kind
            arr["key_off"][n], arr["key_len"][n] = put(o.key.encode())
            arr["uid_off"][n], arr["uid_len"][n] = put(b"u")
            used += pad
            final.append(o.canon)
        n = len(final)
        arr["own_start"][: n + 1] = 0
        arr["dep_start"][: n + 1] = 0
        arr["aux"][: len(aux)] = np.frombuffer(bytes(aux), dtype=np.uint8)
        a.recommit(n, used, len(aux), 0, 0, host.payload_bytes)
        dev = a.download()
        assert [dev.record(i) for i in range(n)] == final
        # the digests the arena holds are those of the new contents, without anybody asking for a re-hash
        h64, ms = ctx.hash_arena(a, A.MODE_HASH64 | A.MODE_CACHED)
        np.testing.assert_array_equal(h64, oracle.h64_arena(dev))
        assert ms < 0.05
        sha, _ = ctx.hash_arena(a, A.MODE_SHA256 | A.MODE_CACHED)
        np.testing.assert_array_equal(sha, oracle.sha256_arena(dev))
        xxh, _ = ctx.hash_arena(a, A.MODE_XXH64 | A.MODE_CACHED)
        assert [int(v) for v in xxh] == [oracle.xxh64(r) for r in final]
        # and a hash-mode scan hashes nothing: the controller still holds its digests, the shard's were carried
        want_codes, want_ents = oracle.scan(host, [dev])
        for m in modes:
            d = ctx.scan(ctrl, [a], m)
            np.testing.assert_array_equal(d.codes(), want_codes)
            np.testing.assert_array_equal(d.entries(), want_ents)
            d.close()
            assert ctx.last_stats().launches == steady[m], (m, ctx.last_stats().launches, steady[m])
    finally:
        a.close()
        ctrl.close()


def test_compaction_reclaims_dead_payload_in_place(ctx):
    """ncd_arena_compact after records were rewritten elsewhere: the live records sit packed at the front on the
    device and in the pinned arrays alike, contents and digests are what they were, scans decide as before, and
    the arena takes value syncs at the new offsets."""
    rng = random.Random(123)
    ctrl_h, shard_hs = random_clusters(55, n_templates=14, n_shards=2)
    arenas = [resident(ctx, h, 1 << 20) for h in [ctrl_h] + shard_hs]
    used = [h.payload_bytes for h in [ctrl_h] + shard_hs]
    try:
        a, arr = arenas[1], arenas[1].host_arrays()
        n = a.counts().nobj
        for rnd in range(3):  # every record of shard 0 grows three times: three dead copies each
            picks = list(range(n))
            for i in picks:
                old = int(arr["rec_len"][i])
                used[1] = rewrite(rng, arr, used[1], i, rand_bytes(rng, old + rng.randint(1, 40)))
            a.sync_records(np.array(picks, dtype=np.uint32), used[1])
        for m in (A.MODE_HASH64, A.MODE_SHA256, A.MODE_XXH64):
            ctx.hash_arena(a, m)
        before = a.download()
        want_codes, want_ents = oracle.scan(arenas[0].download(), [before, arenas[2].download()])
        live = sum((int(before.rec_len[i]) + 15) // 16 * 16 for i in range(n))
        assert used[1] > 2 * live
        assert a.compact() == live
        used[1] = live
        after = a.download()
        assert after.payload_bytes == live
        assert [after.record(i) for i in range(n)] == [before.record(i) for i in range(n)]
        for i in range(n):  # the pinned arrays moved the same way
            off, ln = int(arr["rec_off"][i]), int(arr["rec_len"][i])
            assert off == int(after.rec_off[i]) and arr["payload"][off : off + ln].tobytes() == after.record(i)
        h64, _ = ctx.hash_arena(a, A.MODE_HASH64 | A.MODE_CACHED)
        np.testing.assert_array_equal(h64, oracle.h64_arena(after))
        for mode in (A.MODE_DIRECT, A.MODE_HASH64, A.MODE_SHA256, A.MODE_XXH64):
            d = ctx.scan(arenas[0], arenas[1:], mode)
            np.testing.assert_array_equal(d.codes(), want_codes)
            np.testing.assert_array_equal(d.entries(), want_ents)
            d.close()
        assert a.compact() == live  # nothing left to reclaim
        # value syncs keep working at the new offsets
        i = next(k for k in range(n) if after.rec_len[k] > 0)
        used[1] = rewrite(rng, arr, used[1], i, rand_bytes(rng, int(after.rec_len[i])))
        a.sync_records(np.array([i], dtype=np.uint32), used[1])
        dev = a.download()
        off = int(arr["rec_off"][i])
        assert dev.record(i) == arr["payload"][off : off + int(arr["rec_len"][i])].tobytes()
        d = ctx.scan(arenas[0], arenas[1:], A.MODE_DIRECT)
        np.testing.assert_array_equal(d.codes(), oracle.scan(arenas[0].download(), [dev, arenas[2].download()])[0])
        d.close()
    finally:
        for x in arenas:
            x.close()


def test_compaction_with_records_rewritten_in_place_and_not_synced_yet(ctx):
    """The host layer writes a new value that fits over its record in the pinned arrays as soon as the event
    arrives; a compaction may come before the sync that pulls it.  The host's lengths then differ from the
    device's (shorter, or longer within the record's old room): the slots are laid out for the host's
    lengths, no neighbour is damaged on either side, and after the sync the scan decides like the oracle."""
    rng = random.Random(321)
    ctrl_h, shard_hs = random_clusters(56, n_templates=14, n_shards=2)
    arenas = [resident(ctx, h, 1 << 20) for h in [ctrl_h] + shard_hs]
    used = [h.payload_bytes for h in [ctrl_h] + shard_hs]
    try:
        a, arr = arenas[1], arenas[1].host_arrays()
        n = a.counts().nobj
        for i in range(n):  # dead payload to reclaim: every record grows once
            used[1] = rewrite(rng, arr, used[1], i, rand_bytes(rng, int(arr["rec_len"][i]) + 40))
        a.sync_records(np.arange(n, dtype=np.uint32), used[1])
        # in place, not synced: every third record shrinks, every third grows back into room it has
        touched, expect = [], {}
        for i in range(n):
            ln = int(arr["rec_len"][i])
            if i % 3 == 0 and ln > 20:
                new = rand_bytes(rng, ln - rng.randint(17, min(ln - 1, 200)))
            elif i % 3 == 1:
                new = rand_bytes(rng, ln - ln % 16 + rng.randint(1, 15) if ln % 16 else ln)  # same padded size, other length
            else:
                continue
            off = int(arr["rec_off"][i])
            assert rewrite(rng, arr, used[1], i, new) == used[1] and int(arr["rec_off"][i]) == off  # stayed in place
            touched.append(i)
            expect[i] = new
        untouched = {i: a.download().record(i) for i in range(n) if i not in expect}
        live = sum((int(arr["rec_len"][i]) + 15) // 16 * 16 for i in range(n))
        assert a.compact() == live
        used[1] = live
        for i in range(n):  # pinned arrays: every record whole at its new place
            off, ln = int(arr["rec_off"][i]), int(arr["rec_len"][i])
            assert arr["payload"][off : off + ln].tobytes() == (expect[i] if i in expect else untouched[i]), i
        a.sync_records(np.array(touched, dtype=np.uint32), used[1])
        dev = a.download()
        for i in range(n):
            assert dev.record(i) == (expect[i] if i in expect else untouched[i]), i
        for _ in range(2):
            d = ctx.scan(arenas[0], arenas[1:], A.MODE_DIRECT)
            want_codes, want_ents = oracle.scan(arenas[0].download(), [dev, arenas[2].download()])
            np.testing.assert_array_equal(d.codes(), want_codes)
            np.testing.assert_array_equal(d.entries(), want_ents)
            d.close()
    finally:
        for x in arenas:
            x.close()


def test_committing_several_arenas_at_once_equals_one_by_one(ctx):
    """ncd_arenas_commit (uploads of all arenas enqueued up front, index builds behind them) leaves every arena
    exactly as ncd_arena_commit does, also when the same arenas are committed again and again."""
    from nexus_configuration_controller_b200._ffi import NcdError

    ctrl_h, shard_hs = random_clusters(66, n_templates=20, n_shards=4)
    hosts = [ctrl_h] + shard_hs
    arenas = [Arena(ctx, h.payload_bytes, h.nobj, h.aux_bytes, h.nowner, h.ndep, device_only=False) for h in hosts]
    want_codes, want_ents = oracle.scan(ctrl_h, shard_hs)
    try:
        for a, h in zip(arenas, hosts):
            a.fill_from(h)
        for rep in range(3):
            ctx.commit_many(arenas, [(h.nobj, h.payload_bytes, h.aux_bytes, h.nowner, h.ndep) for h in hosts])
            for mode in (A.MODE_DIRECT, A.MODE_HASH64):
                d = ctx.scan(arenas[0], arenas[1:], mode)
                np.testing.assert_array_equal(d.codes(), want_codes)
                np.testing.assert_array_equal(d.entries(), want_ents)
                d.close()
        for a, h in zip(arenas, hosts):
            dev = a.download()
            assert [dev.record(i) for i in range(h.nobj)] == [h.record(i) for i in range(h.nobj)]
        with pytest.raises(NcdError):  # the same arena twice
            ctx.commit_many([arenas[1], arenas[1]], [(hosts[1].nobj, hosts[1].payload_bytes, hosts[1].aux_bytes, hosts[1].nowner, hosts[1].ndep)] * 2)
        with pytest.raises(NcdError):  # beyond capacity
            ctx.commit_many([arenas[1]], [(hosts[1].nobj + 1, hosts[1].payload_bytes, hosts[1].aux_bytes, hosts[1].nowner, hosts[1].ndep)])
        ctx.commit_many([], [])
    finally:
        for a in arenas:
            a.close()
