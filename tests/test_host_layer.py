"""The host layer (csrc/host: packer, DriftChecker, replay) on two backends.

``fake``  the CPU test double of libncdrift (tests/fake_backend): arenas in plain memory, decisions from
          the oracle.  Runs without a GPU, so the host logic — what goes into the arenas, which upload
          path an event takes, the snapshot and its freshness rule, locking — is covered by the CPU suite.
``gpu``   the real libncdhost.so over libncdrift.so: the same assertions with decisions from the GPU.

The reference's handler tests (tests/golden/reference_cases.json, transcribed from controller_test.go)
are run through both.
"""
import ctypes as C
import os
import random
import subprocess
import threading

import pytest

from nexus_configuration_controller_b200 import arena as A
from nexus_configuration_controller_b200 import host as Hst
from oracle import handlers as H
from tests.reference_fixtures import NS, new_configmap, new_secret, new_template, new_workgroup, with_owners
from tests.test_oracle_golden import CASES

_fake = {}


def fake_lib():
    if "lib" not in _fake:
        from tests.fake_backend.build import build

        _fake["paths"] = build(tsan=False)
        _fake["lib"] = Hst.bind(C.CDLL(_fake["paths"]["lib"]))
    return _fake["lib"]


@pytest.fixture(params=["fake", pytest.param("gpu", marks=pytest.mark.gpu)])
def make_checker(request):
    def make(nshards, mode=0):
        return Hst.DriftChecker(nshards, mode=mode, lib=fake_lib() if request.param == "fake" else None)

    return make


def fill(checker, case):
    for cluster, objs in [(-1, case["controller"])] + list(enumerate(case["shards"])):
        for o in objs:
            checker.add(cluster, o["kind"], o["namespace"], o["name"], o["uid"], bytes.fromhex(o["value_wire"]),
                        [tuple(r) for r in o["owner_refs"]], o["secret_names"], o["configmap_names"])


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_reference_case_through_the_host_layer(make_checker, case):
    """controller_test.go's handler tests: shard actions in order, error or not, and the controller-side
    adoption Updates (expectedControllerUpdateActions, controller_test.go:733-740)."""
    with make_checker(len(case["shards"])) as f:
        fill(f, case)
        f.scan()
        sync = case["sync"]
        plan = f.plan_template if sync["kind"] == "template" else f.plan_workgroup
        nexus, kube, error = plan(sync["namespace"], sync["name"])
        adopt = f.plan_adoption(sync["namespace"], sync["name"]) if sync["kind"] == "template" else []
    exp = case["expect"]
    assert (error is not None) == exp["error"], error
    assert nexus == exp["nexus"]
    assert kube == exp["kube"]
    assert adopt == exp["adopt"]


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

    def apply(self, ops, threads=3):
        """ops: [("put" | "remove", cluster, Obj)] delivered as ONE batch of informer events."""
        keep, evs = [], []
        for op, cluster, o in ops:
            if op == "put":
                self.clusters[cluster][(o.kind, o.namespace, o.name)] = o
                evs.append(Hst.make_event(cluster, o.kind, o.namespace, o.name, o.uid, o.value, o.owner_refs, o.secret_names,
                                          o.configmap_names, keep=keep))
            else:
                self.clusters[cluster].pop((o.kind, o.namespace, o.name), None)
                evs.append(Hst.make_event(cluster, o.kind, o.namespace, o.name, op=Hst.EV_DELETE, keep=keep))
        self.f.apply_events(evs, threads=threads)

    def check(self):
        self.f.scan()
        ctrl = self.clusters[-1]
        shards = [self.clusters[s] for s in range(len(self.clusters) - 1)]
        dirty = []
        for (kind, ns, name), o in ctrl.items():
            adopt = []
            if kind == H.TEMPLATE:
                want = H.template_sync(ctrl, shards, ns, name)
                nexus, kube, err = self.f.plan_template(ns, name)
                adopt = self.f.plan_adoption(ns, name)
                assert adopt == [list(a) for a in H.adopt_references(ctrl, ns, name)], name
            elif kind == H.WORKGROUP:
                want = H.workgroup_sync(ctrl, shards, ns, name)
                nexus, kube, err = self.f.plan_workgroup(ns, name)
            else:
                continue
            assert nexus == [[list(a) for a in sh] for sh in want.nexus_actions], name
            assert kube == [[list(a) for a in sh] for sh in want.kube_actions], name
            assert (err is not None) == (want.error is not None), name
            if any(nexus) or any(kube) or err or adopt:
                dirty.append(("template" if kind == H.TEMPLATE else "workgroup", f"{ns}/{name}"))
        # the sweep feed names exactly the objects whose handler would do something: a shard write, an
        # error, or an adoption Update on the controller cluster
        assert sorted(self.f.sweep_queue()) == sorted(dirty)


def test_resident_arenas_follow_informer_events(make_checker):
    rng = random.Random(5)
    with make_checker(2) as f:
        fleet = Fleet(2, f)
        templates, secrets, configs = [], [], []
        for i in range(6):
            t = new_template(f"algo-{i}", new_secret(f"sec-{i}", None), new_configmap(f"cfg-{i}", None), version_tag=f"v1.{i}")
            templates.append(t)
            # the controller's dependents already carry their template's ownerReference, except the last two
            owner = t if i < 4 else None
            secrets.append(new_secret(f"sec-{i}", owner, {"k": bytes(rng.getrandbits(8) for _ in range(50 + 40 * i))}))
            configs.append(new_configmap(f"cfg-{i}", owner, {"k": "v" * (10 + 30 * i)}))
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
        t = f.timing()
        assert t["full_uploads"] == 3 and t["records_synced"] == 0 and t["h2d_bytes"] > 0 and t["events"] > 0

        # UpdateFunc: values change, same keys and owners -> only those records are pushed
        fleet.put(-1, new_secret("sec-1", templates[1], {"k": b"rotated"}))  # shorter: rewritten in place
        fleet.put(-1, new_configmap("cfg-2", templates[2], {"k": "w" * 5000}))  # longer: appended
        fleet.put(0, with_owners(new_secret("sec-1", None, {"k": b"rotated"}), [templates[1]]))  # shard 0 already converged
        fleet.check()
        assert f.upload_counts(-1) == (1, 0, 1) and f.upload_counts(0) == (1, 0, 1) and f.upload_counts(1) == (1, 0, 0)
        t = f.timing()
        assert (t["full_uploads"], t["structural_uploads"], t["records_synced"], t["events"]) == (0, 0, 3, 3)
        # a put that changes nothing (the 30 s resync re-delivers every object) is no event at all
        fleet.put(-1, new_secret("sec-1", templates[1], {"k": b"rotated"}))
        fleet.check()
        assert f.timing()["events"] == 0 and f.upload_counts(-1) == (1, 0, 1)

        # AddFunc / DeleteFunc / re-owning: the small arrays are re-committed, old records stay put
        fleet.put(1, templates[2])  # shard 1 gets the template it lacked
        fleet.put(1, new_secret("sec-2", None))  # ... and an un-owned (rogue) secret of that name
        fleet.remove(0, with_owners(configs[4], [templates[4]]))
        fleet.put(0, with_owners(secrets[5], [templates[0]]))  # owned by another template only
        fleet.check()
        assert f.upload_counts(0) == (1, 1, 1) and f.upload_counts(1) == (1, 1, 0)
        # adoption done on the controller cluster (what the handler's Update would cause): the feed shrinks
        fleet.put(-1, with_owners(secrets[5], [templates[5]]))
        fleet.put(-1, with_owners(configs[5], [templates[5]]))
        fleet.check()

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


def test_a_snapshot_never_outlives_the_caches_it_was_made_from(make_checker):
    """The reference's handlers read live caches (controller.go:766,792,507,516).  A Secret rotated after a
    scan, a template added or removed after it, a shard copy changed after it: every one of them must be
    answered 'stale', never with the old decision, until the next scan."""
    case = next(c for c in CASES if c["name"] == "TwoShards_converged_and_drifted")
    with make_checker(2) as f:
        fill(f, case)
        f.scan()
        assert f.lookup(A.KIND_TEMPLATE, NS, "test", 1, 0) == A.SKIP
        assert f.lookup(A.KIND_TEMPLATE, NS, "test", 1, 1) == A.UPDATE_DATA
        ctrl_secret = next(o for o in case["controller"] if o["kind"] == A.KIND_SECRET)
        owners = [tuple(r) for r in ctrl_secret["owner_refs"]]
        # the same object delivered again (a resync) changes nothing: still fresh
        f.add(-1, A.KIND_SECRET, NS, ctrl_secret["name"], ctrl_secret["uid"], bytes.fromhex(ctrl_secret["value_wire"]), owners)
        assert f.lookup(A.KIND_TEMPLATE, NS, "test", 1, 0) == A.SKIP
        # rotation on the controller cluster
        rotated = new_secret(ctrl_secret["name"], None, {"secret.file": b"rotated-again"})
        f.add(-1, A.KIND_SECRET, NS, rotated.name, "", rotated.value, owners)
        for shard in (0, 1):
            with pytest.raises(Hst.Stale):
                f.lookup(A.KIND_TEMPLATE, NS, "test", 1, shard)
            with pytest.raises(Hst.Stale):  # the template row too: one of its dependents moved
                f.lookup(A.KIND_TEMPLATE, NS, "test", 0, shard)
        with pytest.raises(Hst.Stale):
            f.plan_template(NS, "test")
        f.scan()
        assert f.lookup(A.KIND_TEMPLATE, NS, "test", 1, 0) == A.UPDATE_DATA  # shard 0 was converged on the old value
        # a change on ONE shard stales that shard's answers only
        shard_cm = next(o for o in case["shards"][1] if o["kind"] == A.KIND_CONFIGMAP)
        f.add(1, A.KIND_CONFIGMAP, NS, shard_cm["name"], "", new_configmap("x", None, {"other": "value"}).value,
              [tuple(r) for r in shard_cm["owner_refs"]])
        assert f.lookup(A.KIND_TEMPLATE, NS, "test", 2, 0) == A.SKIP
        with pytest.raises(Hst.Stale):
            f.lookup(A.KIND_TEMPLATE, NS, "test", 2, 1)
        f.scan()
        # a template created after the scan is not 'absent from the cache': it is not in the snapshot yet
        assert f.lookup(A.KIND_TEMPLATE, NS, "newcomer", 0, 0) is None
        t2 = new_template("newcomer", None, None)
        f.add(-1, A.KIND_TEMPLATE, NS, "newcomer", t2.uid, t2.value)
        with pytest.raises(Hst.Stale):
            f.lookup(A.KIND_TEMPLATE, NS, "newcomer", 0, 0)
        with pytest.raises(Hst.Stale):
            f.plan_template(NS, "newcomer")
        f.scan()
        assert f.lookup(A.KIND_TEMPLATE, NS, "newcomer", 0, 0) == A.CREATE
        # removed after the scan: not NOT_FOUND from the live cache mixed with a code from the old snapshot
        assert f.remove(-1, A.KIND_TEMPLATE, NS, "newcomer")
        with pytest.raises(Hst.Stale):
            f.lookup(A.KIND_TEMPLATE, NS, "newcomer", 0, 0)
        f.scan()
        assert f.lookup(A.KIND_TEMPLATE, NS, "newcomer", 0, 0) is None
        # the shard's copy of the TEMPLATE replaced (new UID): ownership of the dependents hangs on it
        shard_t = next(o for o in case["shards"][0] if o["kind"] == A.KIND_TEMPLATE)
        f.add(0, A.KIND_TEMPLATE, NS, "test", "another-uid", bytes.fromhex(shard_t["value_wire"]), [], shard_t["secret_names"],
              shard_t["configmap_names"])
        with pytest.raises(Hst.Stale):
            f.lookup(A.KIND_TEMPLATE, NS, "test", 1, 0)
        f.scan()
        assert f.lookup(A.KIND_TEMPLATE, NS, "test", 1, 0) & A.UPDATE_OWNER


def test_batched_events_equal_single_puts(make_checker):
    case = next(c for c in CASES if c["name"] == "DataAndOwner_both")
    with make_checker(1) as one, make_checker(1) as batch:
        fill(one, case)
        keep, evs = [], []
        for cluster, objs in [(-1, case["controller"])] + list(enumerate(case["shards"])):
            for o in objs:
                evs.append(Hst.make_event(cluster, o["kind"], o["namespace"], o["name"], o["uid"], bytes.fromhex(o["value_wire"]),
                                          [tuple(r) for r in o["owner_refs"]], o["secret_names"], o["configmap_names"], keep=keep))
        # something to delete again, in the same batch
        evs.append(Hst.make_event(0, A.KIND_SECRET, NS, "short-lived", "", new_secret("x", None).value, keep=keep))
        evs.append(Hst.make_event(0, A.KIND_SECRET, NS, "short-lived", op=Hst.EV_DELETE, keep=keep))
        batch.apply_events(evs, threads=3)
        one.scan()
        batch.scan()
        assert batch.plan_template(NS, "test") == one.plan_template(NS, "test")
        assert batch.counts() == one.counts()
        # a malformed value is skipped and reported: every other event of the batch is in
        good = new_workgroup("wg-a", "shard0")
        bad = [Hst.make_event(-1, A.KIND_WORKGROUP, NS, "wg-a", "u", good.value, keep=keep),
               Hst.make_event(-1, A.KIND_WORKGROUP, NS, "wg-b", "u", b"\xff\x00garbage", keep=keep),
               Hst.make_event(-1, A.KIND_WORKGROUP, NS, "wg-c", "u", good.value, keep=keep)]
        with pytest.raises(Hst.HostError, match="event 1"):
            batch.apply_events(bad, threads=2)
        batch.scan()
        assert batch.lookup(A.KIND_WORKGROUP, NS, "wg-a", 0, 0) == A.CREATE
        assert batch.lookup(A.KIND_WORKGROUP, NS, "wg-b", 0, 0) is None
        assert batch.lookup(A.KIND_WORKGROUP, NS, "wg-c", 0, 0) == A.CREATE


def test_lookups_from_threads_while_another_thread_scans(make_checker):
    """Python threads release the GIL inside the ctypes calls, so this really overlaps lookups, puts and
    scans in the host layer (the native stress test below does it harder and under ThreadSanitizer)."""
    case = next(c for c in CASES if c["name"] == "TwoShards_converged_and_drifted")
    errors, stop = [], threading.Event()
    with make_checker(2) as f:
        fill(f, case)
        f.scan()

        def reader():
            while not stop.is_set():
                try:
                    code = f.lookup(A.KIND_TEMPLATE, NS, "test", 2, 1)
                    if code != A.UPDATE_DATA:
                        errors.append(f"configmap on shard 1 answered {code}")
                    f.plan_template(NS, "test")
                except Hst.Stale:
                    pass
                except Exception as e:  # noqa: BLE001
                    errors.append(repr(e))

        threads = [threading.Thread(target=reader) for _ in range(3)]
        for t in threads:
            t.start()
        wg = new_workgroup("wg", "shard0")
        for k in range(30):
            f.add(-1, A.KIND_WORKGROUP, NS, f"wg-{k % 4}", "u", wg.value)  # unrelated to template "test"
            f.scan()
        stop.set()
        for t in threads:
            t.join()
    assert not errors, errors[:3]


def test_host_layer_under_thread_sanitizer():
    """csrc/host + the CPU backend double built with -fsanitize=thread: writers, scanners and readers at
    once; any data race report fails the run (exit code 66), as does any wrong answer (exit code 1)."""
    from tests.fake_backend.build import build

    paths = build(tsan=True)
    env = dict(os.environ, TSAN_OPTIONS="exitcode=66 halt_on_error=0", NCD_FAKE_SCAN_SLEEP_US="300")
    plain = subprocess.run([paths["stress"], "80"], capture_output=True, text=True, timeout=300, env=env)
    assert plain.returncode == 0, plain.stdout + plain.stderr
    tsan = subprocess.run([paths["stress_tsan"], "40"], capture_output=True, text=True, timeout=600, env=env)
    assert tsan.returncode == 0, (tsan.stdout + tsan.stderr)[-4000:]
    assert "0 failures" in tsan.stdout


@pytest.mark.parametrize("seed", [11, 12])
def test_synthetic_event_batches_decide_like_the_oracle_on_the_raw_arenas(make_checker, request, seed):
    """ncdh_batch_from_arrays wraps every record of arena-shaped arrays as a Go value; equality of the wrapped
    values is equality of the records, so after the events went through canonicaliser, packer and scan every
    decision must equal the oracle's on the raw arenas — shared dependents, dangling references, rogue and
    foreign owners, per-shard object order included.  Then churn: new controller values by salt, shards follow."""
    import oracle
    from tests.helpers import random_clusters

    lib = fake_lib() if "fake" in request.node.name else None
    ctrl, shards = random_clusters(seed, n_templates=14, n_shards=3)

    def check(f, ctrl_h, shard_hs):
        want, _ = oracle.scan(ctrl_h, shard_hs)
        row = 0
        for i in range(ctrl_h.nobj):
            kind = int(ctrl_h.kind[i])
            if kind not in (A.KIND_TEMPLATE, A.KIND_WORKGROUP):
                continue
            ns, name = ctrl_h.key(i).split("/", 1)
            ndeps = int(ctrl_h.dep_start[i + 1] - ctrl_h.dep_start[i]) if kind == A.KIND_TEMPLATE else 0
            for dep in range(ndeps + 1):
                for s in range(len(shard_hs)):
                    assert f.lookup(kind, ns, name, dep, s) == want[row, s], (name, dep, s)
                row += 1
        assert row == want.shape[0]

    with make_checker(3) as f:
        batches = [Hst.EventBatch(ctrl, cluster=-1, threads=2, lib=lib)] + [Hst.EventBatch(h, cluster=s, threads=2, lib=lib) for s, h in enumerate(shards)]
        for b in batches:
            f.apply_events(b, threads=2)
        f.scan()
        check(f, ctrl, shards)
        # churn on the controller: a third of its objects get a new value (one byte, same length)
        rng = random.Random(seed)
        idx = sorted(rng.sample(range(ctrl.nobj), ctrl.nobj // 3))
        upd = Hst.EventBatch(ctrl, cluster=-1, obj_idx=idx, salt=7, threads=2, lib=lib)
        f.apply_events(upd)
        f.scan()
        t = f.timing()
        assert t["full_uploads"] == 0 and t["structural_uploads"] == 0
        changed = [i for i in idx if ctrl.rec_len[i] > 0]
        assert t["records_synced"] == len(changed) and t["events"] == len(changed)
        assert t["records_in_place"] == len(changed)  # same length: written where the record already is, no packer copy
        # the oracle sees the same change when the raw records are patched the same way
        # (ncdh_batch_from_arrays: byte h % len of the record ^= 1 + (h >> 32) % 255, h = mix(salt * P ^ fnv(kind, key)))
        M = (1 << 64) - 1

        def mix(x):
            x = (x + 0x9E3779B97F4A7C15) & M
            x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M
            x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M
            return x ^ (x >> 31)

        def patch(h, i, salt):
            key = h.key(i).encode()
            k = 0xCBF29CE484222325 ^ int(h.kind[i])
            for b in key:
                k = ((k ^ b) * 0x100000001B3) & M
            v = mix(((salt * 0x100000001B3) & M) ^ k)
            ln = int(h.rec_len[i])
            if ln:
                h.payload[int(h.rec_off[i]) + v % ln] ^= 1 + (v >> 32) % 255

        import copy

        ctrl2 = copy.deepcopy(ctrl)
        for i in idx:
            patch(ctrl2, i, 7)
        check(f, ctrl2, shards)
        for b in batches + [upd]:
            b.close()


@pytest.mark.parametrize("seed", [21, 22, 23])
def test_host_layer_and_the_sweep_over_parsed_values_agree(make_checker, request, seed):
    """Two statements of the same sweep fed the same events: the host layer (canonical bytes, arenas, scan)
    and the oracle's fleet (parsed value trees, map lookups, reflect.DeepEqual restated) must produce the
    same diff list — compared by count, by dirty objects and by the layout-independent checksum."""
    import oracle
    from tests.helpers import random_clusters

    lib = fake_lib() if "fake" in request.node.name else None
    ctrl, shards = random_clusters(seed, n_templates=16, n_shards=3)
    _, want_ents = oracle.scan(ctrl, shards)
    fleet = oracle.Fleet(3)
    with make_checker(3) as f:
        rng = random.Random(seed)
        for step in range(3):
            if step == 0:
                batches = [Hst.EventBatch(ctrl, cluster=-1, lib=lib)] + [Hst.EventBatch(h, cluster=s, lib=lib) for s, h in enumerate(shards)]
            else:  # churn: some controller objects change, some shard objects change
                batches = [Hst.EventBatch(ctrl, cluster=-1, obj_idx=rng.sample(range(ctrl.nobj), ctrl.nobj // 4), salt=step, lib=lib)]
                batches += [Hst.EventBatch(h, cluster=s, obj_idx=rng.sample(range(h.nobj), h.nobj // 5), salt=10 + step, lib=lib)
                            for s, h in enumerate(shards)]
            for b in batches:
                f.apply_events(b, threads=2)
                fleet.apply(b.events, b.n, threads=2)
            f.scan()
            _, decisions, ndiff, dirty, digest = fleet.sweep(threads=3)
            assert f.counts() == (ndiff, decisions)
            assert f.diff_digest() == (digest, dirty)
            if step == 0:
                assert ndiff == len(want_ents)
            for b in batches:
                b.close()
    fleet.close()


def test_a_failed_scan_costs_a_full_upload_not_the_cached_values():
    """The pinned arenas are the only copy of the values once a record has been laid out.  A Scan that fails
    on the GPU (here: injected into the CPU double) must leave them intact: the next Scan commits the same
    arenas again in full and decides correctly, including for values that changed before the failure."""
    lib = fake_lib()
    case = next(c for c in CASES if c["name"] == "TwoShards_converged_and_drifted")
    with Hst.DriftChecker(2, lib=lib) as f:
        fill(f, case)
        f.scan()
        want = f.plan_template(NS, "test")
        ctrl_cm = next(o for o in case["controller"] if o["kind"] == A.KIND_CONFIGMAP)
        f.add(-1, A.KIND_CONFIGMAP, NS, ctrl_cm["name"], "", new_configmap("x", None, {"brand": "new"}).value,
              [tuple(r) for r in ctrl_cm["owner_refs"]])
        lib.ncd_fake_fail_next_scans(1)
        with pytest.raises(Hst.HostError, match="injected"):
            f.scan()
        with pytest.raises(Hst.Stale):  # the old snapshot is still there, and still honest about its age
            f.plan_template(NS, "test")
        f.scan()
        assert f.upload_counts(-1)[0] == 2 and f.upload_counts(0)[0] == 2  # every cluster committed again in full
        nexus, kube, err = f.plan_template(NS, "test")
        assert nexus == want[0] and err is None
        assert kube[0] == [["update-data", "configmaps", "test-config"]]  # shard 0 held the old value
        assert kube[1] == want[1][1]
        # and the cheap paths work again afterwards
        f.add(-1, A.KIND_CONFIGMAP, NS, ctrl_cm["name"], "", new_configmap("x", None, {"brand": "newer"}).value,
              [tuple(r) for r in ctrl_cm["owner_refs"]])
        f.scan()
        assert f.upload_counts(-1)[0] == 2 and f.upload_counts(-1)[2] >= 1
        assert f.plan_template(NS, "test")[1][0] == [["update-data", "configmaps", "test-config"]]


def test_an_arena_full_of_dead_payload_is_compacted_not_replaced(make_checker):
    """Values that keep growing are appended behind the used payload and leave their old places dead.  When the
    arena runs out of room and most of it is dead, the host layer has it compacted in place (no upload of the
    cluster) instead of building a larger one; decisions stay right before and after."""
    with make_checker(1) as f:
        fleet = Fleet(1, f)
        cfgs = [new_configmap(f"cfg-{i}", None, {"k": "a" * 3000}) for i in range(8)]
        tmpl = [new_template(f"algo-{i}", None, cfgs[i]) for i in range(8)]
        for i in range(8):
            fleet.put(-1, tmpl[i])
            fleet.put(-1, with_owners(cfgs[i], [tmpl[i]]))
            fleet.put(0, tmpl[i])
            fleet.put(0, with_owners(cfgs[i], [tmpl[i]]))
        fleet.check()
        assert f.upload_counts(-1) == (1, 0, 0)
        # one ConfigMap of the controller grows a little, again and again: every version is appended
        grown = 0
        for step in range(60):
            fleet.put(-1, with_owners(new_configmap("cfg-3", None, {"k": "b" * (3001 + 16 * step)}), [tmpl[3]]))
            # meanwhile another one shrinks, written in place by its batch: when the compaction comes, the
            # pinned arrays hold a record whose length the device does not know yet
            fleet.apply([("put", -1, with_owners(new_configmap("cfg-4", None, {"k": "s" * (2900 - 40 * step)}), [tmpl[4]]))])
            fleet.check()
            grown += 1
            if f.timing()["compactions"] > 0:
                break
        assert f.timing()["compactions"] >= 1, "the arena was never compacted"
        assert f.upload_counts(-1)[0] == 1, "a compaction must not cost a full upload"
        # life goes on in the compacted arena: in-place rewrites, appends, structural changes
        fleet.put(-1, with_owners(new_configmap("cfg-5", None, {"k": "short"}), [tmpl[5]]))
        fleet.put(-1, with_owners(new_configmap("cfg-6", None, {"k": "c" * 5000}), [tmpl[6]]))
        fleet.check()
        fleet.remove(0, with_owners(cfgs[2], [tmpl[2]]))
        fleet.check()
        assert grown > 3


@pytest.mark.parametrize("mode", [0, A.MODE_HASH64], ids=["direct", "hash64"])
@pytest.mark.parametrize("seed", [31, 32, 33, 34])
def test_batches_that_rewrite_records_in_place(make_checker, seed, mode):
    """A batch writes a new value that fits its record's place straight into the pinned payload.  Mixed with
    everything else a batch may carry — values that grow (appended by the packer), shrink, stay the same
    (a resync: nothing to upload), several events for one object, deletions and re-additions, owner changes
    (structural uploads, which must also carry the records rewritten in place) — every plan must still equal
    the handlers' on the same objects after every batch."""
    rng = random.Random(seed)
    n = 6
    with make_checker(2, mode) as f:  # hash64: the digests an arena holds must follow every in-place rewrite and early pull
        fleet = Fleet(2, f)
        size = {}

        def cfg(i, cluster, owners):
            ln, fill = size[(cluster, i)]
            return with_owners(new_configmap(f"cfg-{i}", None, {"k": fill * ln}), owners)

        tmpl = [new_template(f"algo-{i}", None, new_configmap(f"cfg-{i}", None, {})) for i in range(n)]
        first = []
        for i in range(n):
            for cluster in (-1, 0, 1):
                size[(cluster, i)] = (400, "a")
                first += [("put", cluster, tmpl[i]), ("put", cluster, cfg(i, cluster, [tmpl[i]]))]
        fleet.apply(first)
        fleet.check()
        in_place = appended = resyncs = 0
        for _ in range(14):
            ops = []
            for _ in range(rng.randint(2, 10)):
                cluster, i = rng.choice([-1, 0, 1]), rng.randrange(n)
                present = (A.KIND_CONFIGMAP, NS, f"cfg-{i}") in fleet.clusters[cluster]
                what = rng.random()
                if what < 0.5 or not present:  # a new value: same length, shorter or longer
                    ln = size[(cluster, i)][0]
                    ln = rng.choice([ln, ln, max(1, ln - rng.randint(1, 300)), ln + rng.randint(1, 900)])
                    size[(cluster, i)] = (ln, rng.choice("bcdefgh"))
                    ops.append(("put", cluster, cfg(i, cluster, [tmpl[i]])))
                elif what < 0.65:  # the informer's resync: the same object again
                    ops.append(("put", cluster, cfg(i, cluster, [tmpl[i]])))
                    resyncs += 1
                elif what < 0.75:  # two events for one object in the same batch
                    size[(cluster, i)] = (size[(cluster, i)][0], "x")
                    ops.append(("put", cluster, cfg(i, cluster, [tmpl[i]])))
                    size[(cluster, i)] = (size[(cluster, i)][0], "y")
                    ops.append(("put", cluster, cfg(i, cluster, [tmpl[i]])))
                elif what < 0.85 and cluster >= 0:  # gone from a shard, possibly back in the same batch
                    ops.append(("remove", cluster, cfg(i, cluster, [])))
                    if rng.random() < 0.5:
                        ops.append(("put", cluster, cfg(i, cluster, [tmpl[i]])))
                else:  # re-owned: a structural change, with a new value of the same length on top
                    size[(cluster, i)] = (size[(cluster, i)][0], "z")
                    ops.append(("put", cluster, cfg(i, cluster, [] if rng.random() < 0.5 else [tmpl[(i + 1) % n]])))
            before = [f.upload_counts(c) for c in (-1, 0, 1)]
            fleet.apply(ops, threads=rng.choice([1, 3]))
            fleet.check()
            after = [f.upload_counts(c) for c in (-1, 0, 1)]
            in_place += f.timing()["records_in_place"]
            appended += sum(a[1] - b[1] for a, b in zip(after, before))
        assert in_place > 0 and appended > 0 and resyncs > 0


def test_a_large_batch_is_pulled_while_it_is_applied(make_checker, request):
    """A batch of thousands of events is applied in chunks; the records one chunk rewrote in place are pulled to
    the device by the puller thread while the next chunk is canonicalised, and the Scan that follows has nothing
    left to upload.  The diff list must be the oracle fleet's for the same events, before and after."""
    import oracle
    from tests.helpers import random_clusters

    lib = fake_lib() if "fake" in request.node.name else None
    ctrl, shards = random_clusters(77, n_templates=3000, n_shards=2, max_len=600)
    assert ctrl.nobj > 3 * 2048  # several chunks
    fleet = oracle.Fleet(2)
    with make_checker(2) as f:
        first = [Hst.EventBatch(ctrl, cluster=-1, lib=lib)] + [Hst.EventBatch(h, cluster=s, lib=lib) for s, h in enumerate(shards)]
        for b in first:
            f.apply_events(b, threads=4)
            fleet.apply(b.events, b.n, threads=2)
        f.scan()
        _, decisions, ndiff, dirty, digest = fleet.sweep(threads=2)
        assert f.counts() == (ndiff, decisions) and f.diff_digest() == (digest, dirty)
        for salt in (3, 4, 5):  # every controller object gets a new value of the same length, three times
            # (the last time only the objects that have bytes to change: a chunk whose every event was written in place
            # is applied in slices on several workers, the others in order on one)
            idx = None if salt < 5 else [i for i in range(ctrl.nobj) if ctrl.rec_len[i] > 0]
            upd = Hst.EventBatch(ctrl, cluster=-1, obj_idx=idx, salt=salt, threads=4, lib=lib)
            f.apply_events(upd, threads=4)
            fleet.apply(upd.events, upd.n, threads=2)
            f.scan()
            t = f.timing()
            changed = sum(1 for i in range(ctrl.nobj) if ctrl.rec_len[i] > 0)
            assert (t["full_uploads"], t["structural_uploads"]) == (0, 0)
            assert t["records_synced"] == changed and t["records_in_place"] == changed and t["events"] == changed
            _, decisions, ndiff, dirty, digest = fleet.sweep(threads=2)
            assert f.counts() == (ndiff, decisions) and f.diff_digest() == (digest, dirty)
            q = f.sweep_queue()  # long enough to be rendered in slices on the worker pool: still every dirty object once, in order
            assert len(q) == dirty > 2048 and len(set(q)) == dirty
            upd.close()
        for b in first:
            b.close()
    fleet.close()


def test_a_failed_early_pull_fails_the_scan_and_costs_a_full_upload():
    """Records a batch wrote in place are pulled to the device by the puller thread while the batch is still being
    applied.  When such a pull fails (here: injected into the CPU double), the records are ahead of the device
    copy and nobody else knows: the next Scan must report it, publish nothing, and the one after must upload every
    cluster in full from the pinned arenas — the only copy of the values — and decide correctly."""
    lib = fake_lib()
    lib.ncd_fake_fail_next_syncs.argtypes = [C.c_int]
    with Hst.DriftChecker(2, lib=lib) as f:
        fleet = Fleet(2, f)
        cfgs = [new_configmap(f"cfg-{i}", None, {"k": "a" * 200}) for i in range(4)]
        tmpl = [new_template(f"algo-{i}", None, cfgs[i]) for i in range(4)]
        fleet.apply([("put", c, o) for c in (-1, 0, 1) for i in range(4) for o in (tmpl[i], with_owners(cfgs[i], [tmpl[i]]))])
        fleet.check()
        before = f.upload_counts(-1)[0]
        lib.ncd_fake_fail_next_syncs(1)
        fleet.apply([("put", -1, with_owners(new_configmap("cfg-2", None, {"k": "b" * 200}), [tmpl[2]]))])  # same length: in place, pulled early
        with pytest.raises(Hst.HostError, match="injected sync failure"):
            f.scan()
        with pytest.raises(Hst.Stale):
            f.plan_template(NS, "algo-2")
        fleet.check()  # scans again: full uploads, and the plans equal the handlers' on the new value
        assert f.upload_counts(-1)[0] == before + 1 and f.upload_counts(0)[0] == before + 1
        assert f.plan_template(NS, "algo-2")[1][0] == [["update-data", "configmaps", "cfg-2"]]
