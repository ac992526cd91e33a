"""reflect.DeepEqual, rule by rule (CPU).

The compare of the hot path is Go's reflect.DeepEqual (controller.go:539,600,725,795).  It lives in the Go
standard library (go 1.24.4, go.mod:3), which is neither under /root/reference nor buildable here, so its rules
are RESTATED: in oracle/go_values.cpp (deep_equal over value trees), and implicitly in the two canonical
encoders (oracle canon, the product's csrc/host/go_canon.cpp) whose byte equality must coincide with it.

One fixture per sentence of the function's documentation.  ``pinned`` names the reference test that exercises
the rule (controller_test.go); rules no reference test reaches are marked "extrapolated": they follow from the
documentation alone.  Every fixture is checked three ways: oracle.deep_equal, equality of the oracle's canonical
bytes, equality of the product's canonical bytes.
"""
import pytest

import oracle
from nexus_configuration_controller_b200 import host
from tests import govalues as G

DOC = {
    "types": "Values of distinct types are never deeply equal.",
    "struct": "Struct values are deeply equal if their corresponding fields, both exported and unexported, are deeply equal.",
    "iface": "Interface values are deeply equal if they hold deeply equal concrete values.",
    "map": "Map values are deeply equal when all of the following are true: they are both nil or both non-nil, they have the "
           "same length, and either they are the same map object or their corresponding keys (matched using Go equality) map "
           "to deeply equal values.",
    "ptr": "Pointer values are deeply equal if they are equal using Go's == operator or if they point to deeply equal values.",
    "slice": "Slice values are deeply equal when all of the following are true: they are both nil or both non-nil, they have the "
             "same length, and either they point to the same initial entry of the same underlying array or their corresponding "
             "elements (up to length) are deeply equal. Note that a non-nil empty slice and a nil slice are not deeply equal.",
    "other": "Other values - numbers, bools, strings, and channels - are deeply equal if they are equal using Go's == operator.",
}


def quantity(value, scale, cached, fmt="DecimalSI"):
    """resource.Quantity{i int64Amount{value, scale}, d infDecAmount{*inf.Dec}, s string, Format}: all fields
    unexported except Format; `s` caches the canonical string and is filled lazily."""
    return G.Struct(G.Struct(G.Int(value, 64), G.Int(scale, 32)), G.Struct(G.Ptr(None)), G.Str(cached), G.Str(fmt))


E = "extrapolated"
RULES = [
    # (rule, pinned-by or E, a, b, deeply equal?)
    ("map", "TestCreatesSharedResources controller_test.go:1013-1091 ({} == {})", G.BytesMap({}), G.BytesMap({}), True),
    ("map", "TestUpdatesTemplateSecretAndConfig :942-1010 ({} != {k: v})", G.BytesMap({}), G.BytesMap({"secret.file": b"updated-secret"}), False),
    ("map", "TestUpdatesTemplateSecretAndConfig :952-955 (map[string]string)", G.StrMap({}), G.StrMap({"new.file": "updated-config"}), False),
    ("ptr", "TestTakesOwnership :1094-1140 (*bool false vs true)", G.Ptr(G.Bool(False)), G.Ptr(G.Bool(True)), False),
    ("slice", "TestUpdatesWorkgroup :1234-1285 ([] vs [toleration])", G.Slice([]), G.Slice([G.Struct(G.Str("key"), G.Str("Exists"))]), False),
    ("map", E, G.StrMap(None), G.StrMap({}), False),  # both nil or both non-nil
    ("map", E, G.StrMap(None), G.StrMap(None), True),
    ("map", E, G.StrMap({"a": "1", "b": "2"}), G.Map([(G.Str("b"), G.Str("2")), (G.Str("a"), G.Str("1"))]), True),  # no order
    ("map", E, G.StrMap({"a": "1"}), G.StrMap({"a": "1", "b": "2"}), False),  # same length
    ("map", E, G.StrMap({"a": "1"}), G.StrMap({"b": "1"}), False),  # keys matched using Go equality
    ("map", E, G.StrMap({"a": "1"}), G.StrMap({"a": "2"}), False),  # map to deeply equal values
    ("map", E, G.BytesMap({"k": b""}), G.BytesMap({"k": None}), False),  # the VALUES follow the slice rule: empty vs nil
    ("slice", E, G.Slice(None), G.Slice([]), False),
    ("slice", E, G.Bytes(None), G.Bytes(b""), False),  # []byte{} and []byte(nil)
    ("slice", E, G.Slice([G.Str("a"), G.Str("b")]), G.Slice([G.Str("b"), G.Str("a")]), False),  # corresponding elements
    ("slice", E, G.Bytes(b"same-length-a"), G.Bytes(b"same-length-b"), False),
    ("slice", E, G.Slice([G.Int32(1)]), G.Slice([G.Int32(1)]), True),
    ("ptr", E, G.Ptr(None), G.Ptr(None), True),
    ("ptr", E, G.Ptr(None), G.Ptr(G.Int32(0)), False),  # nil vs pointer to the zero value (*int32 DeadlineSeconds)
    ("ptr", E, G.Ptr(G.Struct(G.Str("x"))), G.Ptr(G.Struct(G.Str("x"))), True),  # different addresses, equal pointees
    ("struct", E, G.Struct(G.Str("a"), G.Int32(1)), G.Struct(G.Str("a"), G.Int32(2)), False),
    ("struct", E, G.Struct(G.Str("a"), G.Ptr(G.Struct(G.Slice(None)))), G.Struct(G.Str("a"), G.Ptr(G.Struct(G.Slice(None)))), True),
    # unexported fields count: a Quantity whose cached string has been filled is not DeepEqual to one where it has not
    ("struct", E, quantity(1, 0, "1"), quantity(1, 0, ""), False),
    ("struct", E, quantity(1, 0, "1"), quantity(1, 0, "1"), True),
    ("struct", E, quantity(1000, -3, ""), quantity(1, 0, ""), False),  # numerically equal, represented differently
    ("iface", E, G.Iface(None), G.Iface(None), True),
    ("iface", E, G.Iface(None), G.Iface("int", G.Int(1)), False),
    ("iface", E, G.Iface("int32", G.Int32(1)), G.Iface("myint", G.Int32(1)), False),  # dynamic types differ
    ("iface", E, G.Iface("string", G.Str("v")), G.Iface("string", G.Str("v")), True),
    ("types", E, G.Int(1, 32), G.Int(1, 64), False),
    ("types", E, G.Str("a"), G.Bytes(b"a"), False),
    ("other", E, G.Str("abc"), G.Str("abcd"), False),
    ("other", E, G.Bool(True), G.Bool(True), True),
    ("other", E, G.Int(-1, 64), G.Int(-1, 64), True),
    ("other", E, G.Float(0.0), G.Float(-0.0), True),  # == on floats: +0 == -0 (no float is reachable from the compared types)
]


@pytest.mark.parametrize("k", range(len(RULES)), ids=[f"{r[0]}-{i}" for i, r in enumerate(RULES)])
def test_rule(k):
    rule, pinned, a, b, want = RULES[k]
    assert rule in DOC
    wa, wb = a.wire(), b.wire()
    assert oracle.deep_equal(wa, wb) == want, (DOC[rule], pinned)
    assert oracle.deep_equal(wb, wa) == want  # symmetric
    assert (oracle.canon(wa) == oracle.canon(wb)) == want
    assert (host.canon(wa) == host.canon(wb)) == want
    assert host.canon(wa) == oracle.canon(wa) and host.canon(wb) == oracle.canon(wb)


def test_nan_is_not_deep_equal_to_itself():
    """'Other values ... are deeply equal if they are equal using Go's == operator': NaN != NaN, so no canonical
    encoding can stand for it; both encoders refuse it (no float is reachable from the four compared types)."""
    nan = G.Float(float("nan")).wire()
    assert not oracle.deep_equal(nan, nan)
    with pytest.raises(ValueError):
        oracle.canon(nan)
    with pytest.raises(host.HostError):
        host.canon(nan)


def test_every_documented_rule_has_a_fixture_and_the_pinned_ones_are_named():
    assert {r[0] for r in RULES} == set(DOC)
    pinned = [r[1] for r in RULES if r[1] != E]
    assert len(pinned) == 5 and all("controller_test.go" in p or ":" in p for p in pinned)
