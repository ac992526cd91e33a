"""The reference's sync handlers restated on plain objects — TEST INFRASTRUCTURE.

Follows controller.go line by line for everything that faces a SHARD: templateSyncHandler
(:761-845), syncSecretsToShard (:504-564), syncConfigMapsToShard (:566-626), isMissingOwnership
(:484-502), adoptReferences' existence check (:647-676), workgroupSyncHandler (:697-756).  The
listers are dicts, reflect.DeepEqual is oracle.deep_equal on the value trees (no canonical
encoding, no arenas, no GPU).  Output: the ordered shard write actions the reference's tests
assert on (controller_test.go:588-602 checkActions), and whether the handler returned an error.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

from . import deep_equal

TEMPLATE, WORKGROUP, SECRET, CONFIGMAP = 0, 1, 2, 3
RESOURCE = {TEMPLATE: "NexusAlgorithmTemplates", WORKGROUP: "NexusAlgorithmWorkgroups", SECRET: "secrets", CONFIGMAP: "configmaps"}


@dataclass
class Obj:
    kind: int
    namespace: str
    name: str
    uid: str
    value: bytes  # transport form (tests/govalues.py) of the compared value: .Spec or .Data
    owner_refs: List[Tuple[str, str]] = field(default_factory=list)  # (Kind, UID)
    secret_names: List[str] = field(default_factory=list)  # templates: GetSecretNames()
    configmap_names: List[str] = field(default_factory=list)  # templates: GetConfigMapNames()


Cluster = Dict[Tuple[int, str, str], Obj]  # what the listers of one cluster hold


def cluster(objs: List[Obj]) -> Cluster:
    return {(o.kind, o.namespace, o.name): o for o in objs}


@dataclass
class Result:
    nexus_actions: List[List[Tuple[str, str, str]]]  # per shard: (verb, resource, name), templates / workgroups
    kube_actions: List[List[Tuple[str, str, str]]]  # per shard: secrets / configmaps
    error: Optional[str]


def is_missing_ownership(obj: Obj, owner_uid: Optional[str]):
    """controller.go:484-502.  owner_uid None = the shard template was created in this pass, so
    its UID is one no existing ownerReference can hold."""
    if len(obj.owner_refs) > 0:  # :486
        for kind, uid in obj.owner_refs:  # :489
            if kind == "NexusAlgorithmTemplate" and owner_uid is not None and uid == owner_uid:  # :490
                return False, None
    else:
        return False, f"Resource {obj.name!r} already exists and is not managed by any Algorithm"  # :494-498
    return True, None  # :501


def _sync_dependents(kind: int, names: List[str], ns: str, ctrl: Cluster, shard: Cluster, shard_tmpl_uid, kube: list):
    """syncSecretsToShard :504-564 / syncConfigMapsToShard :566-626."""
    for name in names:  # :505 / :567
        dep = ctrl.get((kind, ns, name))  # :507 / :569
        if dep is None:
            return f"missing {RESOURCE[kind]} {name}"  # :509-514 / :571-576
        sdep = shard.get((kind, ns, dep.name))  # :516 / :578
        if sdep is None:
            kube.append(("create", RESOURCE[kind], name))  # :518-520 / :580-582
            continue  # the created copy has the owner and the data: nothing else fires
        missing, err = is_missing_ownership(sdep, shard_tmpl_uid)  # :529 / :591
        if err:
            return err  # :531-535
        if not deep_equal(dep.value, sdep.value):  # :539 / :600
            kube.append(("update-data", RESOURCE[kind], name))  # :541 / :602
        if missing:
            kube.append(("update-owner", RESOURCE[kind], name))  # :552 / :614
    return None


def template_sync(ctrl: Cluster, shards: List[Cluster], ns: str, name: str) -> Result:
    res = Result([[] for _ in shards], [[] for _ in shards], None)
    t = ctrl.get((TEMPLATE, ns, name))  # :766
    if t is None:
        return res  # :769-772, returns nil
    # adoptReferences (:783 -> :647-676): a referenced object missing on the controller fails the
    # handler before any shard is touched
    for kind, names in ((SECRET, t.secret_names), (CONFIGMAP, t.configmap_names)):
        for n in names:
            if (kind, ns, n) not in ctrl:
                res.error = f"missing {RESOURCE[kind]} {n}"
                return res
    for s, shard in enumerate(shards):  # :790
        st = shard.get((TEMPLATE, ns, name))  # :792
        if st is not None and not deep_equal(st.value, t.value):  # :795
            res.nexus_actions[s].append(("update", RESOURCE[TEMPLATE], name))  # :797
        uid = st.uid if st is not None else None
        if st is None:  # :805
            res.nexus_actions[s].append(("create", RESOURCE[TEMPLATE], name))  # :807
        err = _sync_dependents(SECRET, t.secret_names, ns, ctrl, shard, uid, res.kube_actions[s])  # :817
        if err is None:
            err = _sync_dependents(CONFIGMAP, t.configmap_names, ns, ctrl, shard, uid, res.kube_actions[s])  # :825
        if err is not None:
            res.error = err  # fail fast: later shards are not visited (:819-822, :827-830)
            return res
    return res


def workgroup_sync(ctrl: Cluster, shards: List[Cluster], ns: str, name: str) -> Result:
    res = Result([[] for _ in shards], [[] for _ in shards], None)
    w = ctrl.get((WORKGROUP, ns, name))  # :702
    if w is None:
        return res  # :705-708
    for s, shard in enumerate(shards):  # :720
        sw = shard.get((WORKGROUP, ns, name))  # :722
        if sw is not None and not deep_equal(w.value, sw.value):  # :725
            res.nexus_actions[s].append(("update", RESOURCE[WORKGROUP], name))  # :727
        if sw is None:  # :735
            res.nexus_actions[s].append(("create", RESOURCE[WORKGROUP], name))  # :737
    return res


def adopt_references(ctrl: Cluster, ns: str, name: str) -> List[Tuple[str, str, str]]:
    """adoptReferences (controller.go:647-695): the Updates issued on the CONTROLLER cluster, in order — one
    per referenced Secret, then ConfigMap, that no ownerReference ties to the template yet (isOwnedBy,
    :637-645: UID only).  Stops at the first reference the controller does not hold (:649-653, :672-676)."""
    t = ctrl.get((TEMPLATE, ns, name))
    out: List[Tuple[str, str, str]] = []
    if t is None:
        return out
    for kind, names in ((SECRET, t.secret_names), (CONFIGMAP, t.configmap_names)):
        for n in names:
            dep = ctrl.get((kind, ns, n))
            if dep is None:
                return out
            if not any(uid == t.uid for _, uid in dep.owner_refs):  # :639
                out.append(("update-owner", RESOURCE[kind], n))
    return out
