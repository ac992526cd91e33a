"""bench_steady.py (the steady-state legs of bench.py) at toy size on the CPU test double of the backend:
the GPU arm's host layer and the CPU arm must agree on every step's diff list at every churn level, the
bytes moved must follow the churn, and the shards really catch up one step later."""
import random

import numpy as np

import bench_steady as BS
from nexus_configuration_controller_b200 import arena as A
from tests.helpers import rand_bytes
from tests.test_host_layer import fake_lib


def toy_fleet(groups, S, seed=3):
    rng = random.Random(seed)
    ctrl_objs, per_shard = [], [[] for _ in range(S)]
    for g in range(groups):
        t, s, c = f"nexus/algo-{g}", f"nexus/sec-{g}", f"nexus/cm-{g}"
        vals = [rand_bytes(rng, rng.randint(40, 3000)) for _ in range(3)]
        deps = [(A.KIND_SECRET, s), (A.KIND_CONFIGMAP, c)]
        ctrl_objs += [A.CachedObject(A.KIND_TEMPLATE, t, vals[0], uid=f"c-{g}", deps=deps),
                      A.CachedObject(A.KIND_SECRET, s, vals[1], owners=[(True, f"c-{g}")]),
                      A.CachedObject(A.KIND_CONFIGMAP, c, vals[2], owners=[(True, f"c-{g}")])]
        for k in range(S):
            per_shard[k] += [A.CachedObject(A.KIND_TEMPLATE, t, vals[0], uid=f"s{k}-{g}", deps=deps),
                             A.CachedObject(A.KIND_SECRET, s, vals[1], owners=[(True, f"s{k}-{g}")]),
                             A.CachedObject(A.KIND_CONFIGMAP, c, vals[2], owners=[(True, f"s{k}-{g}")])]
    hosts = [A.pack(ctrl_objs)] + [A.pack(o) for o in per_shard]
    for h in hosts[1:]:
        h.payload = np.zeros(16, dtype=np.uint8)  # shard events take their values from the controller's records
    return hosts


def test_both_arms_walk_the_same_steps():
    lib = fake_lib()
    hosts = toy_fleet(40, 3)
    gpu = BS.SteadyGpu(hosts, device=0, threads=2, lib=lib)
    cpu = BS.SteadyCpu(hosts, threads=2, lib=lib)
    assert gpu.ndiff0 == 0 and gpu.decisions == 40 * 3 * 3 and cpu.last[1] == gpu.decisions and cpu.last[2] == 0
    rows = BS.run_levels(gpu, cpu, hosts, [0.05, 0.25, 1.0], lambda c: (2, 1), threads=2, lib=lib)
    n = hosts[0].nobj
    for churn, row in rows.items():
        assert row["identical_to_cpu"], row
        k = max(1, round(n * churn))
        assert row["records_synced"] == k * 4  # the controller's new values + three shards catching up
        assert row["events_per_step"] == k * 4
        # every changed controller object is one decision != SKIP per shard until the shards follow
        assert row["diff_entries"] == row["cpu_diff_entries"] == k * 3
        assert 0 < row["h2d_bytes_per_step"] < 2 * row["value_bytes_per_step"] + 64 * k * 4
        assert row["decisions_per_s"] > 0 and row["cpu_decisions_per_s"] > 0
        assert set(row["split_ms"]) >= {"canon_ms", "pack_ms", "upload_ms", "scan_ms", "other_ms"}
    assert BS.break_even({0.01: 8.0, 0.1: 2.0, 1.0: 1.5})["churn"] is None
    be = BS.break_even({0.01: 8.0, 0.1: 2.0, 1.0: 0.5})
    assert 0.1 < be["churn"] < 1.0
    gpu.close()
    cpu.close()
