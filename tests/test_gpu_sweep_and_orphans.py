"""Rows f-3 and f-4: the sweep feed (which templates / workgroups still need a handler run) and
the reverse join (what a shard holds that the controller does not), against the oracle / plain sets."""
import numpy as np
import pytest

import oracle
from nexus_configuration_controller_b200 import arena as A
from nexus_configuration_controller_b200.host import DriftChecker
from tests.helpers import random_clusters
from tests.test_gpu_reference_cases import fill
from tests.test_oracle_golden import CASES

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_dirty_objects_are_exactly_the_objects_with_entries(ctx, seed):
    ctrl, shards = random_clusters(seed, n_templates=25, n_shards=3)
    _, want_ents = oracle.scan(ctrl, shards)
    ca = ctx.arena_from_host(ctrl)
    sas = [ctx.arena_from_host(h) for h in shards]
    d = ctx.scan(ca, sas, A.MODE_DIRECT)
    got = d.dirty_objects()
    d.close()
    np.testing.assert_array_equal(got, np.unique(want_ents[:, 0]))
    # a converged fleet has an empty feed
    d = ctx.scan(ca, [], A.MODE_DIRECT)
    assert len(d.dirty_objects()) == 0
    d.close()
    for a in sas + [ca]:
        a.close()


def test_sweep_queue_names_only_what_drifted():
    case = next(c for c in CASES if c["name"] == "TwoShards_converged_and_drifted")
    with DriftChecker(2) as f:
        fill(f, case)
        f.scan()
        assert f.sweep_queue() == [("template", "default/test")]
        # bring shard 1 in line: nothing left to enqueue
        for o in case["controller"]:
            if o["kind"] in (A.KIND_SECRET, A.KIND_CONFIGMAP):
                f.add(1, o["kind"], o["namespace"], o["name"], "", bytes.fromhex(o["value_wire"]), [("NexusAlgorithmTemplate", "test")])
        f.scan()
        assert f.sweep_queue() == []


@pytest.mark.parametrize("seed", [4, 5])
def test_orphans_are_the_set_difference_of_keys(ctx, seed):
    ctrl, shards = random_clusters(seed, n_templates=15, n_shards=4)
    ckeys = {(int(ctrl.kind[i]), ctrl.key(i)) for i in range(ctrl.nobj)}
    want = [(s, i) for s, sh in enumerate(shards) for i in range(sh.nobj) if (int(sh.kind[i]), sh.key(i)) not in ckeys]
    assert want  # every shard carries an object the controller does not know (helpers.random_clusters)
    ca = ctx.arena_from_host(ctrl)
    sas = [ctx.arena_from_host(h) for h in shards]
    got = ctx.find_orphans(ca, sas)
    np.testing.assert_array_equal(got, np.array(want, dtype=np.uint32).reshape(-1, 2))
    # capacity smaller than the answer: the count is still right, and a retry gets everything
    np.testing.assert_array_equal(ctx.find_orphans(ca, sas, cap=1), got)
    # the controller's own objects are never orphans of itself
    assert len(ctx.find_orphans(ca, [ca])) == 0
    for a in sas + [ca]:
        a.close()
