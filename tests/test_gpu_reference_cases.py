"""The reference's handler tests, run against the GPU-backed DriftChecker.

Each case of tests/golden/reference_cases.json is one test of controller_test.go: the objects go
into the controller / shard cache images the way the Go fixture fills the informer indexers
(:546-576), the checker scans once (canonicalise, pack, ncd_scan on the GPU), and the replay of
the handler must issue exactly the shard actions the Go test expects, in order, and fail where it
expects an error.  The same cases pin the CPU oracle in test_oracle_golden.py.
"""
import pytest

from nexus_configuration_controller_b200 import arena as A
from nexus_configuration_controller_b200.host import DriftChecker
from tests.test_oracle_golden import CASES, run_oracle

pytestmark = pytest.mark.gpu

MODES = [A.MODE_DIRECT, A.MODE_HASH64, A.MODE_SHA256 | A.MODE_CONFIRM]


def fill(checker: DriftChecker, case):
    for cluster, objs in [(-1, case["controller"])] + list(enumerate(case["shards"])):
        for o in objs:
            checker.add(cluster, o["kind"], o["namespace"], o["name"], o["uid"], bytes.fromhex(o["value_wire"]),
                        [tuple(r) for r in o["owner_refs"]], o["secret_names"], o["configmap_names"])


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_reference_case(case, mode):
    with DriftChecker(len(case["shards"]), mode=mode) as f:
        fill(f, case)
        f.scan()
        sync = case["sync"]
        plan = f.plan_template if sync["kind"] == "template" else f.plan_workgroup
        nexus, kube, error = plan(sync["namespace"], sync["name"])
    exp = case["expect"]
    assert (error is not None) == exp["error"], error
    assert nexus == exp["nexus"]
    assert kube == exp["kube"]
    # and the oracle's restatement of the handler says the same
    res = run_oracle(case)
    assert nexus == [[list(a) for a in sh] for sh in res.nexus_actions]
    assert kube == [[list(a) for a in sh] for sh in res.kube_actions]


def test_lookup_mirrors_the_call_sites():
    """DriftChecker.Lookup at the sites of controller.go:795 / :539 / :600 (TestUpdatesTemplateSecretAndConfig)."""
    case = next(c for c in CASES if c["name"] == "TestUpdatesTemplateSecretAndConfig")
    with DriftChecker(1) as f:
        fill(f, case)
        f.scan()
        assert f.lookup(A.KIND_TEMPLATE, "default", "test", 0, 0) == A.SKIP  # Spec DeepEqual
        assert f.lookup(A.KIND_TEMPLATE, "default", "test", 1, 0) == A.UPDATE_DATA  # secret .Data differs
        assert f.lookup(A.KIND_TEMPLATE, "default", "test", 2, 0) == A.UPDATE_DATA  # configmap .Data differs
        assert f.lookup(A.KIND_TEMPLATE, "default", "nope", 0, 0) is None  # not in the controller cache
        ndiff, decisions = f.counts()
        assert (ndiff, decisions) == (2, 3)
        # the shard catches up (what UpdateSecret / UpdateConfigMap would do), next scan is clean
        for o in case["controller"]:
            if o["kind"] in (A.KIND_SECRET, A.KIND_CONFIGMAP):
                f.add(0, o["kind"], o["namespace"], o["name"], "", bytes.fromhex(o["value_wire"]), [("NexusAlgorithmTemplate", "test")])
        f.scan()
        assert f.plan_template("default", "test") == ([[]], [[]], None)
        # deleting the shard's template: everything hangs off a template that has to be re-created
        assert f.remove(0, A.KIND_TEMPLATE, "default", "test")
        f.scan()
        nexus, kube, error = f.plan_template("default", "test")
        assert nexus == [[["create", "NexusAlgorithmTemplates", "test"]]] and error is None
        assert kube == [[["update-owner", "secrets", "test-secret"], ["update-owner", "configmaps", "test-config"]]]
