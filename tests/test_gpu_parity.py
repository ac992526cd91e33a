"""GPU parity: libncdrift.so through its C ABI against the CPU oracle, bit for bit.

Bar: integer / byte work, so every decision code and every diff-list entry must be identical.
"""
import ctypes as C

import numpy as np
import pytest

import oracle
from nexus_configuration_controller_b200 import arena as A
from nexus_configuration_controller_b200._ffi import SynthCfg
from tests.helpers import random_clusters, rand_bytes
from tests import synth_spec

pytestmark = pytest.mark.gpu

MODES = [
    A.MODE_DIRECT,
    A.MODE_HASH64,
    A.MODE_HASH64 | A.MODE_CONFIRM,
    A.MODE_SHA256,
    A.MODE_SHA256 | A.MODE_CONFIRM,
    A.MODE_XXH64,
    A.MODE_XXH64 | A.MODE_CONFIRM,
]


def gpu_scan(ctx, ctrl_host, shard_hosts, mode):
    ca = ctx.arena_from_host(ctrl_host)
    sas = [ctx.arena_from_host(h) for h in shard_hosts]
    d = ctx.scan(ca, sas, mode)
    codes, ents = d.codes().copy(), d.entries()
    st = ctx.last_stats()
    d.close()
    for a in sas + [ca]:
        a.close()
    return codes, ents, st


@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5])
@pytest.mark.parametrize("mode", MODES)
def test_random_clusters_match_oracle(ctx, seed, mode):
    ctrl, shards = random_clusters(seed, n_templates=10 + 3 * seed, n_shards=1 + seed % 4)
    want_codes, want_ents = oracle.scan(ctrl, shards)
    codes, ents, st = gpu_scan(ctx, ctrl, shards, mode)
    assert codes.shape == want_codes.shape
    np.testing.assert_array_equal(codes, want_codes)
    np.testing.assert_array_equal(ents, want_ents)
    assert st.ndiff == len(want_ents) and st.pairs == want_codes.size
    # every code must occur somewhere in the seeds, or the test is not exercising the path
    assert set(np.unique(want_codes)) <= {0, 1, 2, 4, 6, 8, 16}


def test_all_codes_are_exercised():
    seen = set()
    for seed in [1, 2, 3, 4, 5]:
        ctrl, shards = random_clusters(seed, n_templates=10 + 3 * seed, n_shards=1 + seed % 4)
        seen |= set(np.unique(oracle.scan(ctrl, shards)[0]).tolist())
    assert seen == {0, 1, 2, 4, 6, 8, 16}


def test_many_shards_more_than_a_warp(ctx):
    ctrl, shards = random_clusters(11, n_templates=6, n_shards=37, max_len=600)
    want_codes, want_ents = oracle.scan(ctrl, shards)
    for mode in (A.MODE_DIRECT, A.MODE_HASH64):
        codes, ents, _ = gpu_scan(ctx, ctrl, shards, mode)
        np.testing.assert_array_equal(codes, want_codes)
        np.testing.assert_array_equal(ents, want_ents)


def test_empty_inputs(ctx):
    empty = A.pack([])
    one = A.pack([A.CachedObject(A.KIND_WORKGROUP, "default/wg", b"\x01\x02\x03", uid="u")])
    # no controller objects: nothing to decide
    codes, ents, st = gpu_scan(ctx, empty, [one], A.MODE_DIRECT)
    assert codes.size == 0 and len(ents) == 0 and st.ndiff == 0
    # no shards
    codes, ents, _ = gpu_scan(ctx, one, [], A.MODE_DIRECT)
    assert codes.size == 0 and len(ents) == 0
    # empty shard cache: everything is created (controller.go:735-738)
    codes, ents, _ = gpu_scan(ctx, one, [empty, one], A.MODE_DIRECT)
    np.testing.assert_array_equal(codes, [[A.CREATE, A.SKIP]])
    np.testing.assert_array_equal(ents, [[0, 0, 0, A.CREATE]])


def test_tail_bytes_beyond_length_are_ignored(ctx):
    """Padding after a record is not part of the value: garbage there must not read as drift."""
    data = bytes(range(1, 38))  # 37 bytes: 5 bytes into the third vector
    ctrl = A.pack([A.CachedObject(A.KIND_WORKGROUP, "default/wg", data)])
    shard = A.pack([A.CachedObject(A.KIND_WORKGROUP, "default/wg", data)])
    shard.payload[37:48] = 0xEE
    for mode in MODES:
        codes, _, _ = gpu_scan(ctx, ctrl, [shard], mode)
        assert codes[0, 0] == A.SKIP, mode
    shard.payload[36] ^= 0x80  # last real byte
    for mode in MODES:
        codes, _, _ = gpu_scan(ctx, ctrl, [shard], mode)
        assert codes[0, 0] == A.UPDATE_DATA, mode


@pytest.mark.parametrize("n", [1, 15, 16, 17, 63, 64, 65, 4095, 4096, 4097, 8192, 100003, 1 << 20])
def test_every_single_byte_position_class_is_seen(ctx, n):
    """Same length, one byte differs: first, last, tile edges (the hard case of cfg 3)."""
    import random

    rng = random.Random(n)
    base = rand_bytes(rng, n)
    positions = sorted({0, n - 1, n // 2, min(n - 1, 4095), min(n - 1, 4096), max(0, n - 17)})
    ctrl = A.pack([A.CachedObject(A.KIND_WORKGROUP, "default/wg", base)])
    shards = []
    for p in positions:
        b = bytearray(base)
        b[p] ^= 0x01
        shards.append(A.pack([A.CachedObject(A.KIND_WORKGROUP, "default/wg", bytes(b))]))
    shards.append(A.pack([A.CachedObject(A.KIND_WORKGROUP, "default/wg", base)]))
    for mode in (A.MODE_DIRECT, A.MODE_HASH64, A.MODE_SHA256, A.MODE_XXH64):
        codes, _, _ = gpu_scan(ctx, ctrl, shards, mode)
        assert codes[0].tolist() == [A.UPDATE_DATA] * len(positions) + [A.SKIP], (mode, n)


def test_digests_match_oracle(ctx):
    import random

    rng = random.Random(99)
    lens = [0, 1, 15, 16, 17, 55, 56, 57, 63, 64, 65, 119, 120, 127, 128, 4095, 4096, 4097, 8192, 12289, 70000]
    objs = [A.CachedObject(A.KIND_SECRET, f"default/s{i}", rand_bytes(rng, n)) for i, n in enumerate(lens)]
    host = A.pack(objs)
    a = ctx.arena_from_host(host)
    h64, _ = ctx.hash_arena(a, A.MODE_HASH64)
    sha, _ = ctx.hash_arena(a, A.MODE_SHA256)
    a.close()
    np.testing.assert_array_equal(h64, oracle.h64_arena(host))
    np.testing.assert_array_equal(sha, oracle.sha256_arena(host))
    import hashlib

    for i, o in enumerate(objs):
        assert bytes(sha[i]) == hashlib.sha256(o.canon).digest()


def test_xxh64_is_the_standard_function(ctx):
    """NCD_MODE_XXH64 digests are XXH64(seed 0): the published known answers, then every tail length
    0..71 (all 8/4/1-byte tail combinations, with and without full stripes) and some long records."""
    import random

    rng = random.Random(64)
    kats = {
        b"": 0xEF46DB3751D8E999,
        b"a": 0xD24EC4F1A98C6E5B,
        b"abc": 0x44BC2CF5AD770999,
        b"Nobody inspects the spammish repetition": 0xFBCEA83C8A378BF1,
    }
    datas = list(kats) + [rand_bytes(rng, n) for n in list(range(72)) + [95, 96, 97, 4095, 4096, 4097, 12289, 70000]]
    objs = [A.CachedObject(A.KIND_SECRET, f"default/s{i}", d) for i, d in enumerate(datas)]
    host = A.pack(objs)
    a = ctx.arena_from_host(host)
    got, _ = ctx.hash_arena(a, A.MODE_XXH64)
    a.close()
    assert got.dtype == np.uint64 and got.shape == (len(datas),)
    for i, d in enumerate(datas):
        assert int(got[i]) == oracle.xxh64(d), (i, len(d))
    for i, want in enumerate(kats.values()):
        assert int(got[i]) == want


def small_synth(**kw) -> SynthCfg:
    c = SynthCfg()
    c.groups, c.n_secrets, c.n_configmaps = 300, 2, 1
    c.tmpl_len_lo, c.tmpl_len_hi = 100, 700
    c.dep_len_lo, c.dep_len_hi = 1, 9000
    c.seed, c.drift_seed = 0xC0FFEE, 0xD21F7
    c.p_flip_ppm, c.p_missing_ppm, c.p_len_ppm = 60000, 40000, 30000
    c.p_owner_ppm, c.p_rogue_ppm = 50000, 30000
    c.part_mod, c.part_rem, c.shuffle = 0, 0, 1
    for k, v in kw.items():
        setattr(c, k, v)
    return c


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("shuffle", [0, 1])
def test_synthetic_arenas_match_oracle_and_spec(ctx, mode, shuffle):
    cfg = small_synth(shuffle=shuffle)
    S = 5
    ca = ctx.synth_arena(cfg, -1)
    sas = [ctx.synth_arena(cfg, s) for s in range(S)]
    d = ctx.scan(ca, sas, mode)
    codes, ents = d.codes().copy(), d.entries()
    d.close()
    ch, shs = ca.download(), [s.download() for s in sas]
    for a in sas + [ca]:
        a.close()
    want_codes, want_ents = oracle.scan(ch, shs)
    np.testing.assert_array_equal(codes, want_codes)
    np.testing.assert_array_equal(ents, want_ents)
    # and the codes are the ones the generator's drift function was asked to inject
    spec = synth_spec.expected_codes(cfg, S)
    np.testing.assert_array_equal(codes, spec)
    assert set(np.unique(codes).tolist()) == {0, 1, 2, 4, 6, 8}


def test_partition_by_key_hash_covers_everything_once(ctx):
    """The multi-GPU layout: each key-hash partition scanned alone gives exactly its share."""
    cfg = small_synth(groups=500)
    S = 3
    whole = synth_spec.expected_codes(cfg, S)
    per = 1 + cfg.n_secrets + cfg.n_configmaps
    kept_total = 0
    G = 4
    for rank in range(G):
        pc = small_synth(groups=500, part_mod=G, part_rem=rank)
        ca = ctx.synth_arena(pc, -1)
        sas = [ctx.synth_arena(pc, s) for s in range(S)]
        d = ctx.scan(ca, sas, A.MODE_DIRECT)
        codes = d.codes().copy()
        d.close()
        for a in sas + [ca]:
            a.close()
        groups = synth_spec.kept_groups(pc)
        kept_total += len(groups)
        want = np.concatenate([whole[g * per : (g + 1) * per] for g in groups]) if len(groups) else whole[:0]
        np.testing.assert_array_equal(codes, want)
    assert kept_total == 500


@pytest.mark.parametrize("variant", [0, 1, 2, 8, 16, "16-unfused"])
def test_every_compare_variant_matches_oracle(variant, monkeypatch):
    """The data-movement shapes of the byte compare (register-staged per shard, grouped, TMA
    bulk-copy pipeline, eight lanes per tile) are picked per scan from the record shapes; forced
    here one by one on records of every size."""
    from nexus_configuration_controller_b200.drift import Context

    if variant == "16-unfused":  # k_join + k_compare_small instead of the fused k_scan_small
        variant = 16
        monkeypatch.setenv("NCD_FUSE_SMALL", "0")
    monkeypatch.setenv("NCD_COMPARE_VARIANT", str(variant))
    with Context(0) as c2:
        for seed, nshards in ((7, 3), (8, 9), (9, 17)):
            ctrl, shards = random_clusters(seed, n_templates=12, n_shards=nshards, max_len=20000)
            want_codes, want_ents = oracle.scan(ctrl, shards)
            for mode in (A.MODE_DIRECT, A.MODE_HASH64 | A.MODE_CONFIRM):
                codes, ents, _ = gpu_scan(c2, ctrl, shards, mode)
                np.testing.assert_array_equal(codes, want_codes)
                np.testing.assert_array_equal(ents, want_ents)
        cfg = small_synth(groups=2000, dep_len_lo=1, dep_len_hi=30000)
        ca = c2.synth_arena(cfg, -1)
        sas = [c2.synth_arena(cfg, s) for s in range(5)]
        d = c2.scan(ca, sas, A.MODE_DIRECT)
        np.testing.assert_array_equal(d.codes(), synth_spec.expected_codes_np(cfg, 5))
        d.close()
        for a in sas + [ca]:
            a.close()


@pytest.mark.parametrize("variant", [0, 2, 8, 16, "16-unfused"])
def test_one_shard_committed_again_between_scans(variant, monkeypatch):
    """The answers of the join are cached per (controller, shard) pair.  When ONE shard is committed again
    (other objects, other numbering) its cache is void while the other shards' caches still hold: the next
    scan must probe that shard and read the caches of the rest — in every kernel that reads them."""
    from nexus_configuration_controller_b200.drift import Arena, Context

    if variant == "16-unfused":
        variant = 16
        monkeypatch.setenv("NCD_FUSE_SMALL", "0")
    monkeypatch.setenv("NCD_COMPARE_VARIANT", str(variant))
    with Context(0) as c2:
        for nshards in (2, 3, 11):
            ctrl, shards = random_clusters(41, n_templates=14, n_shards=nshards, max_len=3000)
            _, others = random_clusters(42, n_templates=14, n_shards=nshards, max_len=3000)
            arenas = [c2.arena_from_host(ctrl)]
            for s in range(nshards):
                h, o = shards[s], others[s]
                a = Arena(c2, max(h.payload_bytes, o.payload_bytes) + 64, max(h.nobj, o.nobj), max(h.aux_bytes, o.aux_bytes) + 64,
                          max(h.nowner, o.nowner) + 1, max(h.ndep, o.ndep) + 1, device_only=False)
                a.fill_from(h)
                a.commit(h.nobj, h.payload_bytes, h.aux_bytes, h.nowner, h.ndep)
                arenas.append(a)
            now = list(shards)
            for step in range(4):
                for _ in range(2):  # the second scan of each state reads what the first one cached
                    d = c2.scan(arenas[0], arenas[1:], A.MODE_DIRECT)
                    want_codes, want_ents = oracle.scan(ctrl, now)
                    np.testing.assert_array_equal(d.codes(), want_codes)
                    np.testing.assert_array_equal(d.entries(), want_ents)
                    d.close()
                s = (step * 2 + 1) % nshards  # one shard gets the other fleet's objects, or its own back
                h = others[s] if now[s] is shards[s] else shards[s]
                arenas[1 + s].fill_from(h)
                arenas[1 + s].commit(h.nobj, h.payload_bytes, h.aux_bytes, h.nowner, h.ndep)
                now[s] = h
            for a in arenas:
                a.close()


def test_scan_into_a_caller_buffer_matches_the_diff_object(ctx):
    """ncd_scan_into: same diff list, written on the device into a caller-owned buffer (the
    all-gather send buffer), no host round trip; too small a buffer is reported, not overrun."""
    import torch

    from nexus_configuration_controller_b200 import multigpu as M

    ctrl, shards = random_clusters(4, n_templates=20, n_shards=4)
    want_codes, want_ents = oracle.scan(ctrl, shards)
    ca = ctx.arena_from_host(ctrl)
    sas = [ctx.arena_from_host(h) for h in shards]
    cap = len(want_ents) + 5
    buf = torch.full((M.buffer_words(cap) + 64,), -1, dtype=torch.int32, device="cuda:0")
    for mode in (A.MODE_DIRECT, A.MODE_HASH64):
        ctx.scan_into(ca, sas, mode, buf.data_ptr(), cap)
        ctx.sync()
        host = buf.cpu().numpy().view(np.uint32)
        (ns, rows, got), = M.unpack_gathered(host[: M.buffer_words(cap)], 1, cap)
        assert (ns, rows) == (len(shards), want_codes.shape[0]) and host[3] == 0
        np.testing.assert_array_equal(got, want_ents)
        assert (host[M.buffer_words(cap):] == 0xFFFFFFFF).all()  # nothing written past the buffer
        st = ctx.stats_back(0)
        assert st.ndiff == len(want_ents) and st.pairs == want_codes.size and st.total_ms > 0
    # a buffer that is too small: header says so, entries up to the capacity are the first ones
    small = 3
    buf.fill_(-1)
    ctx.scan_into(ca, sas, A.MODE_DIRECT, buf.data_ptr(), small)
    ctx.sync()
    host = buf.cpu().numpy().view(np.uint32)
    assert host[0] == len(want_ents) and host[3] == 1
    np.testing.assert_array_equal(host[4 : 4 + 4 * small].reshape(small, 4), want_ents[:small])
    assert (host[4 + 4 * small :] == 0xFFFFFFFF).all()
    for a in sas + [ca]:
        a.close()


def long_key_clusters(seed: int, n_shards: int = 3):
    """Keys longer than the 36 bytes an object line holds, many sharing everything but the last
    bytes, keys of 35 / 36 / 37 bytes with a common prefix, and the Kubernetes maximum
    (63-byte namespace + '/' + 253-byte name)."""
    import random

    rng = random.Random(seed)
    ns_max, name_max = "n" * 63, "m" * 249
    names = (
        [f"{ns_max}/{name_max}{i:04d}" for i in range(40)]                 # 317 bytes, differ in the last 4
        + ["default/" + "a" * k for k in (26, 27, 28, 29, 30)]             # 34 .. 38 bytes
        + ["default/" + "b" * 28 + c for c in "XYZ"]                       # 37 bytes, differ at byte 36
        + ["default/" + "c" * 27 + c for c in "XYZ"]                       # 36 bytes, differ at byte 35
        + [f"team-{i}/" + "shared-long-object-name-with-a-common-prefix-" + "z" * (i % 7) for i in range(30)]
    )
    assert len(set(names)) == len(names)
    secrets = [A.CachedObject(A.KIND_SECRET, k, rand_bytes(rng, rng.randint(1, 3000)), uid=f"c-{k}") for k in names]
    cms = [A.CachedObject(A.KIND_CONFIGMAP, k, rand_bytes(rng, rng.randint(1, 3000)), uid=f"c-{k}") for k in names[::2]]
    templates = []
    for t, k in enumerate(names):
        deps = [(A.KIND_SECRET, d) for d in rng.sample(names, 2)] + [(A.KIND_CONFIGMAP, d) for d in rng.sample(names, 2)]
        templates.append(A.CachedObject(A.KIND_TEMPLATE, k, rand_bytes(rng, rng.randint(1, 5000)), uid=f"tuid-{t}", deps=deps))
    ctrl = templates + secrets + cms
    shards = []
    for s in range(n_shards):
        objs = []
        for o in ctrl:
            roll = rng.random()
            if roll < 0.2:
                continue  # absent
            data = o.canon
            if roll < 0.4:
                b = bytearray(data)
                b[rng.randrange(len(b))] ^= 0x40
                data = bytes(b)
            owners = []
            if o.kind != A.KIND_TEMPLATE:
                refs = [t for t in templates if (o.kind, o.key) in t.deps]
                owners = [(True, f"s{s}-{t.uid}") for t in refs if rng.random() < 0.8]
            objs.append(A.CachedObject(o.kind, o.key, data, uid=f"s{s}-{o.uid}", owners=owners))
        # a near miss of a controller key on the shard only: same first 316 bytes
        objs.append(A.CachedObject(A.KIND_SECRET, f"{ns_max}/{name_max}9999", b"orphan", uid=f"s{s}-orphan"))
        rng.shuffle(objs)
        shards.append(A.pack(objs))
    return A.pack(ctrl), shards


@pytest.mark.parametrize("seed", [1, 2])
def test_long_keys_and_shared_prefixes(ctx, seed):
    ctrl, shards = long_key_clusters(seed)
    want_codes, want_ents = oracle.scan(ctrl, shards)
    assert {0, 1, 2} <= set(np.unique(want_codes).tolist())
    for mode in (A.MODE_DIRECT, A.MODE_HASH64):
        codes, ents, _ = gpu_scan(ctx, ctrl, shards, mode)
        np.testing.assert_array_equal(codes, want_codes)
        np.testing.assert_array_equal(ents, want_ents)
    # the orphan search walks the same keys from the shard side
    ckeys = {(int(ctrl.kind[i]), ctrl.key(i)) for i in range(ctrl.nobj)}
    want = [(s, i) for s, sh in enumerate(shards) for i in range(sh.nobj) if (int(sh.kind[i]), sh.key(i)) not in ckeys]
    assert len(want) == len(shards)
    ca = ctx.arena_from_host(ctrl)
    sas = [ctx.arena_from_host(h) for h in shards]
    np.testing.assert_array_equal(ctx.find_orphans(ca, sas), np.array(want, dtype=np.uint32).reshape(-1, 2))
    for a in sas + [ca]:
        a.close()


def test_many_small_objects_fill_the_hash_table(ctx):
    """60 000 keys per cache: long probe sequences and slots whose stored hash halves collide."""
    import random

    rng = random.Random(5)
    n = 60000
    keys = [f"ns{i % 97}/object-{i}" for i in range(n)]
    datas = [rand_bytes(rng, 1 + (i * 7) % 48) for i in range(n)]
    ctrl = A.pack([A.CachedObject(A.KIND_WORKGROUP, k, d, uid="u") for k, d in zip(keys, datas)])
    order = list(range(n))
    rng.shuffle(order)
    drift = set(rng.sample(range(n), 500))
    gone = set(rng.sample(range(n), 500)) - drift
    shard = A.pack(
        [A.CachedObject(A.KIND_WORKGROUP, keys[i], (datas[i][:-1] + bytes([datas[i][-1] ^ 1])) if i in drift else datas[i], uid="v")
         for i in order if i not in gone]
    )
    want_codes, want_ents = oracle.scan(ctrl, [shard])
    assert (want_codes != 0).sum() == len(drift) + len(gone)
    codes, ents, _ = gpu_scan(ctx, ctrl, [shard], A.MODE_DIRECT)
    np.testing.assert_array_equal(codes, want_codes)
    np.testing.assert_array_equal(ents, want_ents)


def test_small_records_take_the_sub_warp_compare(ctx):
    """Records of a few hundred bytes: the scan picks the fused join + compare kernel (k_scan_small, variant
    32) by itself in direct mode, k_compare_small (16) behind the hash path's confirm pass, and decides like
    the oracle, tails and one-byte differences included — on the first scan (probing) and on the second
    (join answered from the cache the first one filled)."""
    import random

    rng = random.Random(31)
    n = 3000
    datas = [rand_bytes(rng, rng.choice([1, 15, 16, 17, 100, 255, 256, 257, 511, 512])) for _ in range(n)]
    ctrl = A.pack([A.CachedObject(A.KIND_WORKGROUP, f"default/wg-{i}", d, uid="u") for i, d in enumerate(datas)])
    shards = []
    for s in range(9):  # more than one group of eight shards
        objs = []
        for i, d in enumerate(datas):
            roll = rng.random()
            if roll < 0.1:
                continue
            if roll < 0.3:
                b = bytearray(d)
                b[rng.choice([0, len(b) - 1, len(b) // 2])] ^= 1 << rng.randrange(8)
                d = bytes(b)
            elif roll < 0.35:
                d = d + b"x"
            objs.append(A.CachedObject(A.KIND_WORKGROUP, f"default/wg-{i}", d, uid="v"))
        rng.shuffle(objs)
        shards.append(A.pack(objs))
    want_codes, want_ents = oracle.scan(ctrl, shards)
    codes, ents, st = gpu_scan(ctx, ctrl, shards, A.MODE_DIRECT)
    assert st.compare_variant == 32
    np.testing.assert_array_equal(codes, want_codes)
    np.testing.assert_array_equal(ents, want_ents)
    ca = ctx.arena_from_host(ctrl)
    sas = [ctx.arena_from_host(h) for h in shards]
    for _ in range(2):  # the second scan reads the join cache
        d = ctx.scan(ca, sas, A.MODE_DIRECT)
        np.testing.assert_array_equal(d.codes(), want_codes)
        np.testing.assert_array_equal(d.entries(), want_ents)
        d.close()
    for a in sas + [ca]:
        a.close()
    codes, _, st = gpu_scan(ctx, ctrl, shards, A.MODE_HASH64 | A.MODE_CONFIRM)
    assert st.compare_variant == 16
    np.testing.assert_array_equal(codes, want_codes)


@pytest.mark.parametrize("lpo", [4, 8, 16, 32])
@pytest.mark.parametrize("nshards", [3, 8, 9])
def test_fused_small_scan_with_every_lane_grouping(lpo, nshards, monkeypatch):
    """k_scan_small gives an object as many lanes as its records have 16-byte vectors (4 to 32, picked from the
    mean record length); every grouping must decide every record length alike — shorter than one vector per
    lane, exactly one, several rounds — for fewer shards than a group, exactly one group, and more."""
    import random

    from nexus_configuration_controller_b200.drift import Context

    monkeypatch.setenv("NCD_SS_LPO", str(lpo))
    monkeypatch.setenv("NCD_COMPARE_VARIANT", "16")
    rng = random.Random(100 + lpo + nshards)
    n = 1500
    datas = [rand_bytes(rng, rng.choice([0, 1, 16, 63, 64, 65, 127, 128, 255, 256, 257, 511, 512, 513, 1500, 5000])) for _ in range(n)]
    ctrl = A.pack([A.CachedObject(A.KIND_WORKGROUP, f"default/wg-{i}", d, uid="u") for i, d in enumerate(datas)])
    shards = []
    for s in range(nshards):
        objs = []
        for i, d in enumerate(datas):
            roll = rng.random()
            if roll < 0.1:
                continue
            if roll < 0.3 and d:
                b = bytearray(d)
                b[rng.choice([0, len(b) - 1, len(b) // 2])] ^= 1 << rng.randrange(8)
                d = bytes(b)
            elif roll < 0.35:
                d = d + b"x"
            objs.append(A.CachedObject(A.KIND_WORKGROUP, f"default/wg-{i}", d, uid="v"))
        rng.shuffle(objs)
        shards.append(A.pack(objs))
    want_codes, want_ents = oracle.scan(ctrl, shards)
    with Context(0) as c2:
        ca = c2.arena_from_host(ctrl)
        sas = [c2.arena_from_host(h) for h in shards]
        for _ in range(2):  # probing, then from the join cache
            d = c2.scan(ca, sas, A.MODE_DIRECT)
            assert c2.last_stats().compare_variant == 32
            np.testing.assert_array_equal(d.codes(), want_codes)
            np.testing.assert_array_equal(d.entries(), want_ents)
            d.close()
        for a in sas + [ca]:
            a.close()


def test_the_fused_scan_covers_kilobyte_records_but_not_a_skewed_mix(ctx):
    """Records of about a kilobyte still take the fused join + compare kernel (measured ahead of the tile kernels
    up to ~1.5 KB); a fleet of small records with one very long one does not — a handful of lanes would walk that
    record long after everything else is done — and takes the tile kernels, which spread it."""
    import random

    rng = random.Random(77)

    def fleet(lengths):
        datas = [rand_bytes(rng, n) for n in lengths]
        ctrl = A.pack([A.CachedObject(A.KIND_WORKGROUP, f"default/wg-{i}", d, uid="u") for i, d in enumerate(datas)])
        shards = []
        for s in range(3):
            objs = []
            for i, d in enumerate(datas):
                roll = rng.random()
                if roll < 0.1:
                    continue
                if roll < 0.3:
                    b = bytearray(d)
                    b[rng.randrange(len(b))] ^= 1 << rng.randrange(8)
                    d = bytes(b)
                objs.append(A.CachedObject(A.KIND_WORKGROUP, f"default/wg-{i}", d, uid="v"))
            rng.shuffle(objs)
            shards.append(A.pack(objs))
        return ctrl, shards

    for lengths, want_variant in (([rng.randint(900, 1500) for _ in range(1500)], 32),
                                  ([rng.randint(40, 300) for _ in range(1500)] + [100_000], 2)):
        ctrl, shards = fleet(lengths)
        want_codes, want_ents = oracle.scan(ctrl, shards)
        codes, ents, st = gpu_scan(ctx, ctrl, shards, A.MODE_DIRECT)
        assert st.compare_variant == want_variant
        np.testing.assert_array_equal(codes, want_codes)
        np.testing.assert_array_equal(ents, want_ents)
