"""The product's canonicaliser (libncdhost.so) against the oracle's DeepEqual and encoder (CPU)."""
import json
import os
import random

import pytest

import oracle
from nexus_configuration_controller_b200 import host
from tests import govalues as G
from tests.test_oracle_golden import CASES, random_value, shuffled_maps


def test_product_canon_equal_iff_oracle_deep_equal():
    rng = random.Random(77)
    vals = [random_value(rng) for _ in range(500)]
    vals += [shuffled_maps(v, rng) for v in vals[:150]]
    enc = [host.canon(v.wire()) for v in vals]
    for i in range(len(vals)):
        for j in range(i, min(len(vals), i + 50)):
            assert (enc[i] == enc[j]) == oracle.deep_equal(vals[i].wire(), vals[j].wire())
    for i in range(150):
        assert enc[i] == enc[500 + i]
    # two independent statements of the encoding agree byte for byte
    for v, e in zip(vals, enc):
        assert e == oracle.canon(v.wire())


def test_reference_fixture_values_round_trip():
    seen = 0
    for case in CASES:
        for o in case["controller"] + [x for sh in case["shards"] for x in sh]:
            w = bytes.fromhex(o["value_wire"])
            assert host.canon(w) == oracle.canon(w)
            seen += 1
    assert seen > 30


def test_canon_rejects_malformed_and_nan():
    with pytest.raises(host.HostError):
        host.canon(b"\x03\x05\x00\x00\x00ab")  # string shorter than its length
    with pytest.raises(host.HostError):
        host.canon(G.Float(float("nan")).wire())
    with pytest.raises(host.HostError):
        host.canon(G.Str("a").wire() + b"\x00")  # trailing bytes
    with pytest.raises(host.HostError):
        host.canon(G.Map([(G.Str("a"), G.Str("1")), (G.Str("a"), G.Str("2"))]).wire())
    assert host.canon(G.Float(-0.0).wire()) == host.canon(G.Float(0.0).wire())
    assert host.canon(G.StrMap({}).wire()) != host.canon(G.StrMap(None).wire())


def test_header_symbols_exported():
    import ctypes as C
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = re.sub(r"/\*.*?\*/", "", open(os.path.join(root, "include", "ncdhost.h")).read(), flags=re.S)
    names = sorted(set(re.findall(r"\b(ncdh_[a-z0-9_]+)\s*\(", src)))
    lib = C.CDLL(host.LIB_PATH)
    assert all(hasattr(lib, n) for n in names) and sorted(host.PROTOTYPES) == names
