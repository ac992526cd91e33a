"""Seeded random cluster snapshots for the parity tests (host-side, numpy arenas)."""
from __future__ import annotations

import random
from typing import List, Tuple

from nexus_configuration_controller_b200.arena import (
    KIND_CONFIGMAP,
    KIND_SECRET,
    KIND_TEMPLATE,
    KIND_WORKGROUP,
    CachedObject,
    HostArena,
    pack,
)

EDGE_LENGTHS = [1, 2, 15, 16, 17, 31, 32, 33, 511, 512, 4095, 4096, 4097, 8191, 8192, 8193, 12288, 70001]


def rand_bytes(rng: random.Random, n: int) -> bytes:
    return rng.getrandbits(8 * n).to_bytes(n, "little") if n else b""


def random_clusters(seed: int, n_templates: int = 12, n_shards: int = 3, max_len: int = 9000) -> Tuple[HostArena, List[HostArena]]:
    """A controller cache and n_shards shard caches exercising every decision code.

    Templates reference 0-3 secrets and 0-3 configmaps out of a shared pool (so dependents are
    shared between templates, controller_test.go:1013-1091), some references dangle
    (ERR_MISSING_REF).  Every shard copy is independently: absent, equal, same length with one
    byte changed, different length; dependents additionally: owned by the right template, by a
    foreign one, by nobody (rogue).  Object order differs per shard.
    """
    rng = random.Random(seed)
    ns = "default"

    def length() -> int:
        return rng.choice(EDGE_LENGTHS) if rng.random() < 0.4 else rng.randint(1, max_len)

    pool_s = [f"{ns}/secret-{i}" for i in range(max(2, n_templates // 2))]
    pool_c = [f"{ns}/config-{i}" for i in range(max(2, n_templates // 2))]
    ctrl: List[CachedObject] = []
    dep_data = {}
    for key in pool_s:
        dep_data[(KIND_SECRET, key)] = rand_bytes(rng, length())
    for key in pool_c:
        dep_data[(KIND_CONFIGMAP, key)] = rand_bytes(rng, length())
    templates = []
    for t in range(n_templates):
        deps = [(KIND_SECRET, k) for k in rng.sample(pool_s, rng.randint(0, min(3, len(pool_s))))]
        if rng.random() < 0.15:
            deps.append((KIND_SECRET, f"{ns}/dangling-secret-{t}"))
        deps += [(KIND_CONFIGMAP, k) for k in rng.sample(pool_c, rng.randint(0, min(3, len(pool_c))))]
        if rng.random() < 0.15:
            deps.append((KIND_CONFIGMAP, f"{ns}/dangling-config-{t}"))
        o = CachedObject(KIND_TEMPLATE, f"{ns}/algo-{t}", rand_bytes(rng, length()), uid=f"ctrl-uid-{t}", deps=deps)
        templates.append(o)
    workgroups = [
        CachedObject(KIND_WORKGROUP, f"{ns}/wg-{w}", rand_bytes(rng, length()), uid=f"wg-uid-{w}") for w in range(max(1, n_templates // 4))
    ]
    # a secret and a configmap may share a name: kinds are separate stores
    dep_data[(KIND_SECRET, f"{ns}/same-name")] = rand_bytes(rng, 100)
    dep_data[(KIND_CONFIGMAP, f"{ns}/same-name")] = rand_bytes(rng, 100)
    templates[0].deps = [(KIND_SECRET, f"{ns}/same-name")] + templates[0].deps + [(KIND_CONFIGMAP, f"{ns}/same-name")]
    templates[0].deps.sort(key=lambda d: d[0])  # secrets first, then configmaps
    ctrl = templates + workgroups + [CachedObject(k, key, data, uid=f"cuid-{key}") for (k, key), data in dep_data.items()]
    rng.shuffle(ctrl)

    def mutate(data: bytes) -> bytes:
        r = rng.random()
        if r < 0.55:
            return data
        if r < 0.8:  # same length, one byte changed (first, last or random position)
            b = bytearray(data)
            pos = rng.choice([0, len(b) - 1, rng.randrange(len(b))])
            b[pos] ^= rng.randint(1, 255)
            return bytes(b)
        if r < 0.9:
            return data + rand_bytes(rng, rng.randint(1, 20))
        return data[: max(0, len(data) - rng.randint(1, 5))]

    shards = []
    for s in range(n_shards):
        objs: List[CachedObject] = []
        present_templates = {}
        for o in ctrl:
            if rng.random() < 0.2:
                continue  # absent from this shard
            if o.kind in (KIND_TEMPLATE, KIND_WORKGROUP):
                so = CachedObject(o.kind, o.key, mutate(o.canon), uid=f"s{s}-{o.uid}")
                if o.kind == KIND_TEMPLATE:
                    present_templates[o.key] = so.uid
            else:
                r = rng.random()
                referencing = [t for t in templates if (o.kind, o.key) in t.deps]
                owners: List[Tuple[bool, str]] = []
                if r < 0.15:
                    owners = []  # rogue
                elif r < 0.3:
                    owners = [(True, "someone-else"), (False, f"s{s}-ctrl-uid-0")]  # foreign / wrong kind
                else:
                    owners = [(True, f"s{s}-{t.uid}") for t in referencing if rng.random() < 0.8]
                    if not owners:
                        owners = [(True, "unrelated-template")]
                so = CachedObject(o.kind, o.key, mutate(o.canon), uid=f"s{s}-{o.uid}", owners=owners)
            objs.append(so)
        # objects the controller does not know about are ignored by the reference (no orphan sweep)
        objs.append(CachedObject(KIND_SECRET, f"{ns}/orphan-{s}", rand_bytes(rng, 64), owners=[(True, "x")]))
        rng.shuffle(objs)
        shards.append(pack(objs))
    return pack(ctrl), shards
