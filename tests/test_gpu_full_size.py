"""BASELINE.json's configs at FULL size on one B200, checked through size-independent properties.

The oracle cannot walk 50-110 GB in seconds, so every decision is checked through size-independent
properties, and the oracle itself reads payload on a SAMPLE (drifted groups + a seeded draw of clean
ones, read back from the resident arenas: oracle_on_a_sample).  The properties:
  - injected drift is found exactly: the generator's drift function (restated in
    tests/synth_spec.py) says which (row, shard) decisions must be non-SKIP and with which code;
    the scan must return exactly those — every same-length one-byte change among tens of GB found,
    nothing else reported;
  - the diff list is the codes != SKIP, in (row, shard) order, once (compaction is exact);
  - idempotence: a second scan of the same arenas gives the same list;
  - mode independence: the hash path decides exactly like the byte-compare path.
"""
import numpy as np
import pytest

import bench
from nexus_configuration_controller_b200 import arena as A
from tests import synth_spec

pytestmark = pytest.mark.gpu


def _keys(meta):
    """Every key of an arena image as fixed-width numpy bytes (the generator's keys are at most 22 bytes)."""
    n = meta.nobj
    off = meta.key_off[:n].astype(np.int64)
    k = meta.aux[(off[:, None] + np.arange(24)[None, :]).clip(max=len(meta.aux) - 1)]
    k[np.arange(24)[None, :] >= meta.key_len[:n, None]] = 0
    return np.ascontiguousarray(k).view("S24").ravel()


def _objects(meta, idx, records, with_deps):
    def text(off, ln):
        return meta.aux[int(off) : int(off) + int(ln)].tobytes().decode()

    out = []
    for i, rec in zip(idx, records):
        owners = [(bool(meta.own_is_template[o]), text(meta.own_uid_off[o], meta.own_uid_len[o]))
                  for o in range(int(meta.own_start[i]), int(meta.own_start[i + 1]))]
        deps = [(int(meta.dep_kind[d]), text(meta.dep_key_off[d], meta.dep_key_len[d]))
                for d in range(int(meta.dep_start[i]), int(meta.dep_start[i + 1]))] if with_deps else []
        out.append(A.CachedObject(int(meta.kind[i]), meta.key(i), rec, uid=text(meta.uid_off[i], meta.uid_len[i]), owners=owners, deps=deps))
    return out


def oracle_on_a_sample(ctx, cfg, S, ca, sas, codes, n_drifted, n_clean, seed=5):
    """the oracle reads PAYLOAD at full size too, on a sample — drifted groups and a seeded
    draw of clean ones.  Their records are read back from the resident arenas (every cluster copy, found by
    key in the shard's own object order), laid out as small clusters on the host, decided by the CPU oracle
    (hash-map listers + memcmp) and compared with the rows the full-size scan produced for those groups."""
    import oracle

    per = 1 + cfg.n_secrets + cfg.n_configmaps
    rng = np.random.default_rng(seed)
    drifted = np.unique(np.nonzero(codes)[0] // per)
    clean = np.setdiff1d(np.arange(cfg.groups), drifted, assume_unique=True)
    pick = np.sort(np.concatenate([rng.choice(drifted, size=min(n_drifted, len(drifted)), replace=False),
                                   rng.choice(clean, size=min(n_clean, len(clean)), replace=False)])).astype(np.int64)
    cidx = (pick[:, None] * per + np.arange(per)[None, :]).ravel()
    cmeta = ca.download_meta()
    ckeys = _keys(cmeta)[cidx]
    ctrl_mini = A.pack(_objects(cmeta, cidx, ca.fetch_records(cidx), with_deps=True))
    shard_minis, fetched = [], 0
    for sa in sas:
        smeta = sa.download_meta()
        skeys = _keys(smeta)
        order = np.argsort(skeys, kind="stable")
        at = np.searchsorted(skeys[order], ckeys)
        at[at >= len(order)] = 0
        sidx = order[at]
        sidx = sidx[skeys[sidx] == ckeys]  # objects this shard lacks are simply not in its small cluster
        recs = sa.fetch_records(sidx)
        fetched += sum(len(r) for r in recs)
        shard_minis.append(A.pack(_objects(smeta, sidx, recs, with_deps=False)))
    want_codes, _ = oracle.scan(ctrl_mini, shard_minis, nthreads=8)
    rows = (pick[:, None] * per + np.arange(per)[None, :]).ravel()
    assert want_codes.shape == (len(rows), S)
    bad = np.argwhere(want_codes != codes[rows])
    assert len(bad) == 0, f"oracle sample: {len(bad)} decisions differ, first {bad[:5].tolist()}"
    return len(pick), int(np.count_nonzero(want_codes)), fetched


def run_config(ctx, name, modes, sample=(0, 0)):
    cfg, S = bench.make_cfg(name)
    ca = ctx.synth_arena(cfg, -1)
    sas = [ctx.synth_arena(cfg, s) for s in range(S)]
    want = synth_spec.expected_codes_np(cfg, S)
    want_n = int(np.count_nonzero(want))
    first = None
    try:
        for mode in modes:
            for rep in range(2):
                d = ctx.scan(ca, sas, mode)
                codes, ents = d.codes(), d.entries()
                st = ctx.last_stats()
                d.close()
                assert codes.shape == want.shape
                bad = np.argwhere(codes != want)
                assert len(bad) == 0, f"{name} mode {mode}: {len(bad)} decisions differ, first {bad[:5].tolist()}"
                assert st.ndiff == want_n == len(ents)
                if want_n:
                    rows, shards = np.nonzero(want)
                    per = 1 + cfg.n_secrets + cfg.n_configmaps
                    np.testing.assert_array_equal(ents[:, 0], (rows // per) * per)  # owning template's index
                    np.testing.assert_array_equal(ents[:, 1], rows % per)
                    np.testing.assert_array_equal(ents[:, 2], shards)
                    np.testing.assert_array_equal(ents[:, 3], want[rows, shards])
                if first is None:
                    first = ents
                else:
                    np.testing.assert_array_equal(ents, first)
        if sum(sample):
            groups, found, nbytes = oracle_on_a_sample(ctx, cfg, S, ca, sas, codes, *sample)
            print(f"{name}: oracle walked {groups} sampled groups ({nbytes / 1e6:.0f} MB of shard payload), {found} decisions != SKIP among them")
            assert found > 0 or want_n == 0
        return st, want_n
    finally:
        for a in sas + [ca]:
            a.close()


def test_cfg2_pure_compare_path_no_drift(ctx):
    st, n = run_config(ctx, "cfg2_10k_templates_4KB_x8_nodrift", [A.MODE_DIRECT, A.MODE_HASH64, A.MODE_SHA256, A.MODE_XXH64], sample=(0, 2000))
    assert n == 0 and st.pairs == 240_000 and st.algo_bytes == 783_360_000  # SURVEY.md 8(d) cfg 2


def test_cfg3_100k_configmaps_64KB_16_shards_1pct_drift(ctx):
    st, n = run_config(ctx, "cfg3_100k_configmaps_64KB_x16_1pct", [A.MODE_DIRECT], sample=(400, 200))
    assert st.pairs == 3_200_000 and 25_000 < n < 40_000  # 1 % of 3.2 M decisions
    assert st.algo_bytes > 111_000_000_000


def test_cfg4_1M_mixed_objects_8_shards(ctx):
    st, n = run_config(ctx, "cfg4_1M_mixed_1-16KB_x8", [A.MODE_DIRECT, A.MODE_HASH64, A.MODE_XXH64], sample=(2000, 1000))
    assert st.pairs == 8_000_016 and 8_000 < n < 10_500
    assert 77e9 < st.algo_bytes < 80e9


def test_cfg5_10k_secrets_1MiB_4_shards(ctx):
    st, n = run_config(ctx, "cfg5_10k_secrets_1MB_x4", [A.MODE_DIRECT, A.MODE_HASH64], sample=(0, 60))
    assert n == 0 and st.pairs == 80_000 and st.algo_bytes > 52_428_800_000
