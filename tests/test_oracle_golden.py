"""Pins the oracle: the reference's own handler tests, as transcribed data, must come out right (CPU)."""
import json
import os
import random

import pytest

import oracle
from oracle import handlers as H
from tests import govalues as G

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = json.load(open(os.path.join(HERE, "golden", "reference_cases.json")))["cases"]


def load_obj(j) -> H.Obj:
    return H.Obj(j["kind"], j["namespace"], j["name"], j["uid"], bytes.fromhex(j["value_wire"]),
                 [tuple(r) for r in j["owner_refs"]], list(j["secret_names"]), list(j["configmap_names"]))


def run_oracle(case) -> H.Result:
    ctrl = H.cluster([load_obj(o) for o in case["controller"]])
    shards = [H.cluster([load_obj(o) for o in sh]) for sh in case["shards"]]
    fn = H.template_sync if case["sync"]["kind"] == "template" else H.workgroup_sync
    return fn(ctrl, shards, case["sync"]["namespace"], case["sync"]["name"])


def test_the_fixture_file_is_what_the_script_writes():
    from tests.golden import make_reference_cases as m

    assert m.build()["cases"] == CASES
    pinned = [c["name"] for c in CASES if c["pinned"]]
    assert pinned == [
        "TestCreatesTemplate", "TestDetectsRogue", "TestHandlesNotExistingResource", "TestSkipsInvalidTemplate",
        "TestUpdatesTemplateSecretAndConfig", "TestCreatesSharedResources", "TestTakesOwnership",
        "TestCreatesWorkgroup", "TestUpdatesWorkgroup",
    ]


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_oracle_handlers_reproduce_the_reference_tests(case):
    res = run_oracle(case)
    exp = case["expect"]
    assert (res.error is not None) == exp["error"], res.error
    assert [[list(a) for a in sh] for sh in res.nexus_actions] == exp["nexus"]
    assert [[list(a) for a in sh] for sh in res.kube_actions] == exp["kube"]


# ---------------------------------------------------------------- reflect.DeepEqual rules
def eq(a, b):
    return oracle.deep_equal(a.wire(), b.wire())


def test_deep_equal_rules():
    # pinned by the reference tests: {} == {} (TestCreatesSharedResources), {} != {k:v}
    # (TestUpdatesTemplateSecretAndConfig), *bool deref (TestTakesOwnership), [] != [x] (TestUpdatesWorkgroup)
    assert eq(G.BytesMap({}), G.BytesMap({}))
    assert not eq(G.BytesMap({}), G.BytesMap({"secret.file": b"updated-secret"}))
    assert not eq(G.Ptr(G.Bool(True)), G.Ptr(G.Bool(False)))
    assert eq(G.Ptr(G.Bool(True)), G.Ptr(G.Bool(True)))
    assert not eq(G.Slice([]), G.Slice([G.Str("x")]))
    # extrapolated from the documented rules
    assert not eq(G.StrMap(None), G.StrMap({}))  # nil map vs empty map
    assert not eq(G.Slice(None), G.Slice([]))  # nil slice vs empty slice
    assert not eq(G.Bytes(None), G.Bytes(b""))  # nil []byte vs empty []byte
    assert eq(G.Ptr(None), G.Ptr(None)) and not eq(G.Ptr(None), G.Ptr(G.Int32(0)))
    assert eq(G.Map([(G.Str("a"), G.Str("1")), (G.Str("b"), G.Str("2"))]), G.Map([(G.Str("b"), G.Str("2")), (G.Str("a"), G.Str("1"))]))
    assert not eq(G.StrMap({"a": "1"}), G.StrMap({"a": "2"})) and not eq(G.StrMap({"a": "1"}), G.StrMap({"b": "1"}))
    assert not eq(G.Bytes(b"abc"), G.Bytes(b"abd")) and not eq(G.Str("abc"), G.Str("abcd"))
    assert not eq(G.Int(1, 32), G.Int(1, 64))  # different types are never equal
    assert not eq(G.Str("a"), G.Bytes(b"a"))
    assert eq(G.Float(0.0), G.Float(-0.0)) and not eq(G.Float(float("nan")), G.Float(float("nan")))
    assert eq(G.Iface(None), G.Iface(None)) and not eq(G.Iface(None), G.Iface("int", G.Int(1)))
    assert not eq(G.Iface("int32", G.Int32(1)), G.Iface("myint", G.Int32(1)))
    assert eq(G.Struct(G.Str("a"), G.Ptr(G.Struct(G.Slice(None)))), G.Struct(G.Str("a"), G.Ptr(G.Struct(G.Slice(None)))))


def random_value(rng: random.Random, depth=0) -> G.V:
    kinds = ["bool", "int", "str", "bytes", "nil"] + (["slice", "map", "ptr", "struct"] if depth < 4 else [])
    k = rng.choice(kinds)
    if k == "bool":
        return G.Bool(rng.random() < 0.5)
    if k == "int":
        return G.Int(rng.choice([0, 1, -1, 2**31 - 1, rng.randint(-5, 5)]), rng.choice([32, 64]))
    if k == "str":
        return G.Str(rng.choice(["", "a", "b", "ab", "key", "a\x00"]))
    if k == "bytes":
        return G.Bytes(rng.choice([b"", b"a", b"\x00", b"ab"]))
    if k == "nil":
        return rng.choice([G.Slice(None), G.Map(None), G.Ptr(None), G.Bytes(None)])
    if k == "slice":
        return G.Slice([random_value(rng, depth + 1) for _ in range(rng.randint(0, 3))])
    if k == "ptr":
        return G.Ptr(random_value(rng, depth + 1))
    if k == "struct":
        return G.Struct(*[random_value(rng, depth + 1) for _ in range(rng.randint(0, 3))])
    keys = rng.sample(["", "a", "b", "c", "aa"], rng.randint(0, 4))
    return G.Map([(G.Str(x), random_value(rng, depth + 1)) for x in keys])


def shuffled_maps(v: G.V, rng: random.Random) -> G.V:
    """Same value with every map's entries in another insertion order."""
    if v.t == G.T_MAP:
        ents = [(k, shuffled_maps(x, rng)) for k, x in v.v]
        rng.shuffle(ents)
        return G.V(v.t, ents)
    if v.t in (G.T_SLICE, G.T_STRUCT):
        return G.V(v.t, [shuffled_maps(x, rng) for x in v.v])
    if v.t == G.T_PTR:
        return G.V(v.t, shuffled_maps(v.v, rng))
    return v


def test_canonical_bytes_equal_iff_deep_equal():
    rng = random.Random(2024)
    vals = [random_value(rng) for _ in range(400)]
    vals += [shuffled_maps(v, rng) for v in vals[:100]]
    canon = [oracle.canon(v.wire()) for v in vals]
    n_equal = 0
    for i in range(len(vals)):
        for j in range(i, min(len(vals), i + 60)):
            de = oracle.deep_equal(vals[i].wire(), vals[j].wire())
            assert de == (canon[i] == canon[j]), (vals[i], vals[j])
            n_equal += de
    for i in range(100):  # map order never matters
        assert canon[i] == canon[400 + i] and oracle.deep_equal(vals[i].wire(), vals[400 + i].wire())
    assert n_equal > 500


def test_canon_refuses_nan():
    with pytest.raises(ValueError):
        oracle.canon(G.Float(float("nan")).wire())
    assert oracle.canon(G.Float(0.0).wire()) == oracle.canon(G.Float(-0.0).wire())


def test_deep_equal_walk_timing_helper():
    """The second CPU baseline figure (bench.py cpu_baseline.deep_equal_walk): the timing helper
    really compares, says so when values differ, and the bench wrapper reports sane numbers."""
    import bench
    from tests import govalues as G

    a = G.BytesMap({"token": bytes(range(200)) * 10}).wire()
    b = G.BytesMap({"token": bytes(range(200)) * 9 + bytes(range(199)) + b"\x00"}).wire()
    sec, eq = oracle.deep_equal_timed(a, a, 64, 2, 2)
    assert sec > 0 and eq
    sec, eq = oracle.deep_equal_timed(a, b, 64, 1, 3)
    assert sec > 0 and not eq
    with pytest.raises(ValueError):
        oracle.deep_equal_timed(a[:-3], a, 4, 1, 1)
    r = bench.deep_equal_walk_baseline(512, 2, budget_s=0.2)
    assert r["all_equal"] and r["value"] > 1e4 and r["cores"] == 2 and r["unit"] == "pairs/s"
