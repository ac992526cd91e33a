"""BASELINE.json's configs at FULL size on one B200, checked through size-independent properties.

The oracle cannot walk 50-110 GB in seconds, so these tests do not read payload on the CPU.
What they check instead:
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


def run_config(ctx, name, modes):
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
        return st, want_n
    finally:
        for a in sas + [ca]:
            a.close()


def test_cfg2_pure_compare_path_no_drift(ctx):
    st, n = run_config(ctx, "cfg2_10k_templates_4KB_x8_nodrift", [A.MODE_DIRECT, A.MODE_HASH64, A.MODE_SHA256, A.MODE_XXH64])
    assert n == 0 and st.pairs == 240_000 and st.algo_bytes == 783_360_000  # SURVEY.md 8(d) cfg 2


def test_cfg3_100k_configmaps_64KB_16_shards_1pct_drift(ctx):
    st, n = run_config(ctx, "cfg3_100k_configmaps_64KB_x16_1pct", [A.MODE_DIRECT])
    assert st.pairs == 3_200_000 and 25_000 < n < 40_000  # 1 % of 3.2 M decisions
    assert st.algo_bytes > 111_000_000_000


def test_cfg4_1M_mixed_objects_8_shards(ctx):
    st, n = run_config(ctx, "cfg4_1M_mixed_1-16KB_x8", [A.MODE_DIRECT, A.MODE_HASH64, A.MODE_XXH64])
    assert st.pairs == 8_000_016 and 8_000 < n < 10_500
    assert 77e9 < st.algo_bytes < 80e9


def test_cfg5_10k_secrets_1MiB_4_shards(ctx):
    st, n = run_config(ctx, "cfg5_10k_secrets_1MB_x4", [A.MODE_DIRECT, A.MODE_HASH64])
    assert n == 0 and st.pairs == 80_000 and st.algo_bytes > 52_428_800_000
