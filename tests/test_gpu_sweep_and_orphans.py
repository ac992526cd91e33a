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


@pytest.mark.parametrize("seed", [6, 7])
def test_unadopted_references_match_the_oracle(ctx, seed):
    """ncd_find_unadopted (adoptReferences, controller.go:647-695, as a list) against the CPU restatement: a
    controller cluster whose dependents are owned by their template, by another template, by several, by a
    non-template owner with the same UID (isOwnedBy compares UIDs only), or by nobody; dangling references."""
    import ctypes as C
    import random

    rng = random.Random(seed)
    objs, names = [], []
    for t in range(40):
        deps = []
        for k, kind in ((0, A.KIND_SECRET), (1, A.KIND_CONFIGMAP), (2, A.KIND_SECRET)):
            key = f"default/dep-{t}-{k}"
            deps.append((kind, key))
            how = rng.randrange(6)
            if how == 0:
                continue  # dangling: the controller does not hold it
            owners = {1: [], 2: [(True, f"uid-{t}")], 3: [(True, "uid-other")], 4: [(True, "uid-x"), (False, f"uid-{t}"), (True, "uid-y")],
                      5: [(False, f"uid-{t}")]}[how]
            objs.append(A.CachedObject(kind, key, b"v", owners=owners))
        deps.sort(key=lambda d: d[0])
        objs.append(A.CachedObject(A.KIND_TEMPLATE, f"default/algo-{t}", b"spec", uid=f"uid-{t}", deps=deps))
    rng.shuffle(objs)
    host = A.pack(objs)
    want = oracle.unadopted(host)
    assert len(want) > 10
    a = ctx.arena_from_host(host)
    out = np.zeros((len(want) + 8, 2), dtype=np.uint32)
    n = C.c_uint64()
    from nexus_configuration_controller_b200 import _ffi

    _ffi.check(ctx.lib.ncd_find_unadopted(ctx.handle, a.handle, out.ctypes.data_as(_ffi.u32p), len(out), C.byref(n)))
    np.testing.assert_array_equal(out[: n.value], want)
    # a buffer that is too small: the count is still the whole answer
    _ffi.check(ctx.lib.ncd_find_unadopted(ctx.handle, a.handle, out.ctypes.data_as(_ffi.u32p), 3, C.byref(n)))
    assert n.value == len(want)
    a.close()
