"""Writes tests/golden/reference_cases.json: the reference's handler tests as data.

    python tests/golden/make_reference_cases.py

The reference (Go) cannot be run here, and it has no golden files for this path; what it has are
handler tests whose expectations are ordered lists of client actions (controller_test.go:588-602).
Each case below transcribes one test: the objects it puts in the controller / shard informer
caches (via the builders mirrored in tests/reference_fixtures.py) and the SHARD actions it
expects — copied from the Go test's expect* calls, never computed by this repository's code.
Controller-side actions (status updates, adoption) are outside the drift-check path.

Cases marked pinned=false extend the reference's tests in the direction SURVEY.md §8c lists as
not covered by any reference test; their expectations follow from Go semantics and are flagged
"extrapolated".
"""
from __future__ import annotations

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from tests.reference_fixtures import (  # noqa: E402
    NS, new_configmap, new_secret, new_template, new_workgroup, toleration, with_owners, with_uid,
)

T, W, S, C = "NexusAlgorithmTemplates", "NexusAlgorithmWorkgroups", "secrets", "configmaps"


def obj_json(o):
    return {
        "kind": o.kind, "namespace": o.namespace, "name": o.name, "uid": o.uid, "value_wire": o.value.hex(),
        "owner_refs": [list(r) for r in o.owner_refs], "secret_names": o.secret_names, "configmap_names": o.configmap_names,
    }


def case(name, source, sync, controller, shards, nexus, kube, error, pinned=True, note="", adopt=()):
    """adopt: the Updates adoptReferences issues on the CONTROLLER cluster (controller.go:647-695), as the Go
    tests expect them through expectedControllerUpdateActions (controller_test.go:733-740); empty where a test
    expects none."""
    return {
        "name": name, "source": source, "pinned": pinned, "note": note,
        "sync": {"kind": sync[0], "namespace": NS, "name": sync[1]},
        "controller": [obj_json(o) for o in controller],
        "shards": [[obj_json(o) for o in sh] for sh in shards],
        "expect": {"error": error, "nexus": nexus, "kube": kube, "adopt": [list(a) for a in adopt]},
    }


def build():
    cases = []
    BOTH = [["update-owner", S, "test-secret"], ["update-owner", C, "test-config"]]  # expectedControllerUpdateActions :733-740
    # provisionControllerResources (:747-752)
    secret, config = new_secret("test-secret", None), new_configmap("test-config", None)
    template = new_template("test", secret, config)

    # TestCreatesTemplate :800-843 — empty shard; expectShardActions(..., false) :835-839
    cases.append(case("TestCreatesTemplate", "controller_test.go:800-843", ("template", "test"),
                      [template, secret, config], [[]],
                      nexus=[[["create", T, "test"]]], kube=[[["create", S, "test-secret"], ["create", C, "test-config"]]], error=False,
                      adopt=BOTH))  # :833

    # TestDetectsRogue :846-886 — shard holds the un-owned secret (:863-865); template created, then
    # fail-fast: "no actions expected" for secret/configmap (:875-880); f.run(..., true) expects error
    cases.append(case("TestDetectsRogue", "controller_test.go:846-886", ("template", "test"),
                      [template, secret, config], [[secret]],
                      nexus=[[["create", T, "test"]]], kube=[[]], error=True, adopt=BOTH))  # :882

    # TestHandlesNotExistingResource :889-909 — template not in the controller cache: nil, no actions
    cases.append(case("TestHandlesNotExistingResource", "controller_test.go:889-909", ("template", "test"),
                      [], [[]], nexus=[[]], kube=[[]], error=False))

    # TestSkipsInvalidTemplate :912-939 — referenced secret/configmap absent on the controller: error, no shard action
    cases.append(case("TestSkipsInvalidTemplate", "controller_test.go:912-939", ("template", "test"),
                      [template], [[]], nexus=[[]], kube=[[]], error=True))

    # TestUpdatesTemplateSecretAndConfig :942-1010
    sec_upd = with_owners(new_secret("test-secret", None, {"secret.file": b"updated-secret"}), [template])  # :947-950
    cfg_upd = with_owners(new_configmap("test-config", None, {"new.file": "updated-config"}), [template])  # :952-955
    tmpl_ctrl = new_template("test", sec_upd, cfg_upd)  # :957
    tmpl_shard = new_template("test", secret, config)  # templateOnShard :970 (labels differ, not compared)
    sec_shard, cfg_shard = new_secret("test-secret", tmpl_shard), new_configmap("test-config", tmpl_shard)  # :971,977
    cases.append(case("TestUpdatesTemplateSecretAndConfig", "controller_test.go:942-1010", ("template", "test"),
                      [tmpl_ctrl, sec_upd, cfg_upd], [[tmpl_shard, sec_shard, cfg_shard]],
                      nexus=[[]], kube=[[["update-data", S, "test-secret"], ["update-data", C, "test-config"]]], error=False))

    # TestCreatesSharedResources :1013-1091 — template2 created; dependents exist, owned by template1 only
    t1, t2 = new_template("test1", secret, config), new_template("test2", secret, config)  # :1017,1027
    own1_s, own1_c = with_owners(secret, [t1]), with_owners(config, [t1])  # :1025
    t1_shard = with_uid(t1, "test1")  # :1045
    s_shard1, c_shard1 = with_owners(secret, [t1_shard]), with_owners(config, [t1_shard])  # :1043-1044
    cases.append(case("TestCreatesSharedResources", "controller_test.go:1013-1091", ("template", "test2"),
                      [t1, t2, own1_s, own1_c], [[t1_shard, s_shard1, c_shard1]],
                      nexus=[[["create", T, "test2"]]],  # :1084
                      kube=[[["update-owner", S, "test-secret"], ["update-owner", C, "test-config"]]],  # :1085-1087
                      error=False, adopt=BOTH))  # :1083: template2 is appended to the owners template1 left

    # TestTakesOwnership :1094-1140 — shard template differs in *bool MountDatadogSocket (:1099-1100)
    rogue_t = with_uid(new_template("test", secret, config, mount_datadog_socket=False), "")
    cases.append(case("TestTakesOwnership", "controller_test.go:1094-1140", ("template", "test"),
                      [template, secret, config], [[rogue_t]],
                      nexus=[[["update", T, "test"]]],  # expectShardActions(..., true) :1132-1136
                      kube=[[["create", S, "test-secret"], ["create", C, "test-config"]]], error=False, adopt=BOTH))  # :1131

    # TestCreatesWorkgroup :1193-1231
    wg = new_workgroup("test-workgroup", "shard0")
    cases.append(case("TestCreatesWorkgroup", "controller_test.go:1193-1231", ("workgroup", "test-workgroup"),
                      [wg], [[]], nexus=[[["create", W, "test-workgroup"]]], kube=[[]], error=False))

    # TestUpdatesWorkgroup :1234-1285 — Tolerations [] on the shard vs one element on the controller
    wg_upd = new_workgroup("test-workgroup", "shard0", [toleration("key", "Exists")])  # :1239-1244
    cases.append(case("TestUpdatesWorkgroup", "controller_test.go:1234-1285", ("workgroup", "test-workgroup"),
                      [wg_upd], [[wg]], nexus=[[["update", W, "test-workgroup"]]], kube=[[]], error=False))

    # ---- extrapolated (SURVEY.md 8c "not pinned by any reference test")
    # BASELINE.json configs[0]: 1 template + 1 secret + 1 configmap to TWO shards: shard0 converged,
    # shard1 as in TestUpdatesTemplateSecretAndConfig
    sec_ok, cfg_ok = with_owners(sec_upd, [tmpl_shard]), with_owners(cfg_upd, [tmpl_shard])
    cases.append(case("TwoShards_converged_and_drifted", "extends controller_test.go:942-1010 to 2 shards (BASELINE.json configs[0])",
                      ("template", "test"), [tmpl_ctrl, sec_upd, cfg_upd],
                      [[tmpl_shard, sec_ok, cfg_ok], [tmpl_shard, sec_shard, cfg_shard]],
                      nexus=[[], []], kube=[[], [["update-data", S, "test-secret"], ["update-data", C, "test-config"]]],
                      error=False, pinned=False, note="extrapolated: the reference tests never use more than one shard"))
    # data differs AND owner is foreign: two writes, data first (controller.go:539-560)
    foreign = new_template("other", None, None, uid="other-uid")
    sec_both = with_owners(new_secret("test-secret", None, {"secret.file": b"stale"}), [foreign])
    cases.append(case("DataAndOwner_both", "controller.go:539-560 (no reference test)", ("template", "test"),
                      [tmpl_ctrl, sec_upd, cfg_upd], [[tmpl_shard, sec_both, cfg_ok]],
                      nexus=[[]], kube=[[["update-data", S, "test-secret"], ["update-owner", S, "test-secret"]]],
                      error=False, pinned=False, note="extrapolated"))
    # nil map vs empty map are not DeepEqual (reflect.DeepEqual doc: "both nil or both non-nil")
    from oracle.handlers import SECRET as K_SECRET, Obj
    from tests.govalues import BytesMap

    sec_nil = Obj(K_SECRET, NS, "test-secret", "", BytesMap(None).wire(), owner_refs=[("NexusAlgorithmTemplate", "test")])
    cases.append(case("NilVsEmptyMap", "reflect.DeepEqual map rule (no reference test)", ("template", "test"),
                      [template, secret, config], [[tmpl_shard, sec_nil, with_owners(config, [tmpl_shard])]],
                      nexus=[[]], kube=[[["update-data", S, "test-secret"]]], error=False, pinned=False, note="extrapolated",
                      adopt=BOTH))
    # rogue on the SECOND shard: first shard's actions stand, the handler stops at the rogue (fail-fast :819-822)
    cases.append(case("RogueOnSecondShard_failfast", "controller.go:790-831 fail-fast (no reference test)", ("template", "test"),
                      [template, secret, config], [[], [tmpl_shard, secret], []],
                      nexus=[[["create", T, "test"]], [], []],
                      kube=[[["create", S, "test-secret"], ["create", C, "test-config"]], [], []],
                      error=True, pinned=False, note="extrapolated: third shard is never visited", adopt=BOTH))
    return {"generated_by": "tests/golden/make_reference_cases.py", "cases": cases}


if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_cases.json")
    with open(out, "w") as f:
        json.dump(build(), f, indent=1)
    print("wrote", out)
