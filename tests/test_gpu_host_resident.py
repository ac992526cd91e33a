"""The DriftChecker keeps every cluster's arena resident: value changes, structural changes and
growth beyond capacity each take their own upload path, and after every one of them the replayed
handler actions must equal the oracle's restatement of the handlers on the same objects."""
import random

import pytest

from nexus_configuration_controller_b200 import arena as A
from nexus_configuration_controller_b200.host import DriftChecker
from oracle import handlers as H
from tests import govalues as G
from tests.reference_fixtures import NS, new_configmap, new_secret, new_template, new_workgroup, with_owners

pytestmark = pytest.mark.gpu


class Fleet:
    """Controller + shards as plain object dicts, mirrored into a DriftChecker."""

    def __init__(self, nshards, checker):
        self.clusters = {c: {} for c in range(-1, nshards)}
        self.f = checker

    def put(self, cluster, o: H.Obj):
        self.clusters[cluster][(o.kind, o.namespace, o.name)] = o
        self.f.add(cluster, o.kind, o.namespace, o.name, o.uid, o.value, o.owner_refs, o.secret_names, o.configmap_names)

    def remove(self, cluster, o: H.Obj):
        del self.clusters[cluster][(o.kind, o.namespace, o.name)]
        assert self.f.remove(cluster, o.kind, o.namespace, o.name)

    def check(self):
        self.f.scan()
        ctrl = self.clusters[-1]
        shards = [self.clusters[s] for s in range(len(self.clusters) - 1)]
        dirty = []
        for (kind, ns, name), o in ctrl.items():
            if kind == H.TEMPLATE:
                want = H.template_sync(ctrl, shards, ns, name)
                nexus, kube, err = self.f.plan_template(ns, name)
            elif kind == H.WORKGROUP:
                want = H.workgroup_sync(ctrl, shards, ns, name)
                nexus, kube, err = self.f.plan_workgroup(ns, name)
            else:
                continue
            assert nexus == [[list(a) for a in sh] for sh in want.nexus_actions], name
            assert kube == [[list(a) for a in sh] for sh in want.kube_actions], name
            assert (err is not None) == (want.error is not None), name
            if any(nexus) or any(kube) or err:
                dirty.append(("template" if kind == H.TEMPLATE else "workgroup", f"{ns}/{name}"))
        # the sweep feed names exactly the objects whose handler would do something (a missing
        # reference is an error before any shard is touched, and shows as a decision too)
        assert sorted(self.f.sweep_queue()) == sorted(dirty)


def test_resident_arenas_follow_informer_events():
    rng = random.Random(5)
    with DriftChecker(2) as f:
        fleet = Fleet(2, f)
        secrets = [new_secret(f"sec-{i}", None, {"k": bytes(rng.getrandbits(8) for _ in range(50 + 40 * i))}) for i in range(6)]
        configs = [new_configmap(f"cfg-{i}", None, {"k": "v" * (10 + 30 * i)}) for i in range(6)]
        templates = [new_template(f"algo-{i}", secrets[i], configs[i], version_tag=f"v1.{i}") for i in range(6)]
        for o in secrets + configs + templates + [new_workgroup("wg", "shard0")]:
            fleet.put(-1, o)
        for s in (0, 1):
            for i in range(6):
                if (i + s) % 3 == 0:
                    continue  # this shard lacks template i and its dependents
                fleet.put(s, templates[i])
                fleet.put(s, with_owners(secrets[i], [templates[i]]))
                fleet.put(s, with_owners(configs[i], [templates[i]]))
            fleet.put(s, new_workgroup("wg", "shard0"))
        fleet.check()
        assert f.upload_counts(-1) == (1, 0, 0) and f.upload_counts(0) == (1, 0, 0)

        # UpdateFunc: values change, same keys and owners -> only those records are pushed
        fleet.put(-1, new_secret("sec-1", None, {"k": b"rotated"}))  # shorter: rewritten in place
        fleet.put(-1, new_configmap("cfg-2", None, {"k": "w" * 5000}))  # longer: appended
        fleet.put(0, with_owners(new_secret("sec-1", None, {"k": b"rotated"}), [templates[1]]))  # shard 0 already converged
        fleet.check()
        assert f.upload_counts(-1) == (1, 0, 1) and f.upload_counts(0) == (1, 0, 1) and f.upload_counts(1) == (1, 0, 0)

        # AddFunc / DeleteFunc / re-owning: the small arrays are re-committed, old records stay put
        fleet.put(1, templates[2])  # shard 1 gets the template it lacked
        fleet.put(1, new_secret("sec-2", None))  # ... and an un-owned (rogue) secret of that name
        fleet.remove(0, with_owners(configs[4], [templates[4]]))
        fleet.put(0, with_owners(secrets[5], [templates[0]]))  # owned by another template only
        fleet.check()
        assert f.upload_counts(0) == (1, 1, 1) and f.upload_counts(1) == (1, 1, 0)

        # growth beyond the arena's spare room -> one fresh full upload, then back to cheap syncs
        for i in range(6, 40):
            sec, cfg = new_secret(f"sec-{i}", None, {"k": b"x" * 3000}), new_configmap(f"cfg-{i}", None, {"k": "y" * 3000})
            for o in (sec, cfg, new_template(f"algo-{i}", sec, cfg)):
                fleet.put(-1, o)
        fleet.check()
        assert f.upload_counts(-1)[0] == 2
        fleet.put(-1, new_secret("sec-7", None, {"k": b"z" * 3000}))
        fleet.check()
        full, structural, values = f.upload_counts(-1)
        assert (full, values) == (2, 2)
