"""The object builders of the reference's own tests, transcribed (controller_test.go:230-379).

Each function names the Go function it mirrors.  Only what the drift check reads is modelled:
the compared value (.Spec / .Data) as a Go value tree, name/namespace/UID, ownerReferences and
the names a template references.  Field order inside a struct follows the literal in the Go
test; nexus-core's real declaration order (v1.4.1, not in the tree) is irrelevant to equality as
long as both sides are built the same way.
"""
from __future__ import annotations

from typing import Dict, List, Optional

from oracle.handlers import CONFIGMAP, SECRET, TEMPLATE, WORKGROUP, Obj
from tests.govalues import Bool, BytesMap, Int, Int32, Map, Ptr, Slice, Str, StrMap, Struct, V

NS = "default"  # metav1.NamespaceDefault


def env_from_source(secret_name: Optional[str] = None, configmap_name: Optional[str] = None) -> V:
    """corev1.EnvFromSource{Prefix, ConfigMapRef *ConfigMapEnvSource, SecretRef *SecretEnvSource}."""
    def src(name):  # {LocalObjectReference{Name}, Optional *bool}
        return Ptr(Struct(Struct(Str(name)), Ptr(None))) if name is not None else Ptr(None)

    return Struct(Str(""), src(configmap_name), src(secret_name))


def template_spec(secret_name: Optional[str], configmap_name: Optional[str], mount_datadog_socket: bool = True,
                  version_tag: str = "v1.0.0") -> V:
    """nexusv1.NexusAlgorithmSpec as newTemplate builds it (controller_test.go:292-325)."""
    return Struct(
        Ptr(Struct(Str("test"), Str("test"), Str(version_tag), Str("test"))),  # Container :293-298
        Ptr(Struct(Str("1000m"), Str("2000Mi"), StrMap(None))),  # ComputeResources :299-303 (CustomResources nil)
        Ptr(Struct(Str("test"), Str("test"), Str("NexusAlgorithmWorkgroup"))),  # WorkgroupRef :304-308
        Str("python"),  # Command :309
        Slice([Str("job.py")]),  # Args :310 (cargs, :262,267)
        Ptr(Struct(  # RuntimeEnvironment :311-317
            Slice(None),  # EnvironmentVariables nil
            Slice([env_from_source(secret_name=secret_name), env_from_source(configmap_name=configmap_name)]),  # :261,269-282
            StrMap(None),  # Annotations nil
            Ptr(None),  # DeadlineSeconds nil
            Ptr(None),  # MaximumRetries nil
        )),
        Ptr(Struct(Slice([]), Slice([]))),  # ErrorHandlingBehaviour :318-321: make([]int32, 0) twice
        Ptr(Struct(Ptr(Bool(mount_datadog_socket)))),  # DatadogIntegrationSettings :322-324
    )


def new_template(name: str, secret: Optional[Obj], configmap: Optional[Obj], uid: Optional[str] = None, **spec_kw) -> Obj:
    """newTemplate (controller_test.go:260-333): UID = name (:290)."""
    return Obj(
        TEMPLATE, NS, name, name if uid is None else uid,
        template_spec(secret.name if secret else None, configmap.name if configmap else None, **spec_kw).wire(),
        secret_names=[secret.name] if secret else [],
        configmap_names=[configmap.name] if configmap else [],
    )


def new_secret(name: str, owner: Optional[Obj], data: Optional[Dict[str, bytes]] = None) -> Obj:
    """newSecret (controller_test.go:335-356): Data = make(map[string][]byte), non-nil and empty."""
    refs = [("NexusAlgorithmTemplate", owner.uid)] if owner is not None else []
    return Obj(SECRET, NS, name, "", BytesMap({} if data is None else data).wire(), owner_refs=refs)


def new_configmap(name: str, owner: Optional[Obj], data: Optional[Dict[str, str]] = None) -> Obj:
    """newConfigMap (controller_test.go:358-379): Data = make(map[string]string)."""
    refs = [("NexusAlgorithmTemplate", owner.uid)] if owner is not None else []
    return Obj(CONFIGMAP, NS, name, "", StrMap({} if data is None else data).wire(), owner_refs=refs)


def toleration(key: str, operator: str) -> V:
    """corev1.Toleration{Key, Operator, Value, Effect, TolerationSeconds *int64}."""
    return Struct(Str(key), Str(operator), Str(""), Str(""), Ptr(None))


def workgroup_spec(cluster: str, tolerations: Optional[List[V]] = None) -> V:
    """nexusv1.NexusAlgorithmWorkgroupSpec as newWorkgroup builds it (controller_test.go:244-250)."""
    return Struct(
        Str("test workgroup"),  # Description
        Map([]),  # Capabilities: make(map[string]bool)
        Str(cluster),  # Cluster
        Slice([] if tolerations is None else tolerations),  # Tolerations: []corev1.Toleration{}
        Ptr(Struct(Ptr(None), Ptr(None), Ptr(None))),  # Affinity: &corev1.Affinity{}
    )


def new_workgroup(name: str, cluster: str, tolerations: Optional[List[V]] = None) -> Obj:
    """newWorkgroup (controller_test.go:230-258): UID = name (:242)."""
    return Obj(WORKGROUP, NS, name, name, workgroup_spec(cluster, tolerations).wire())


def with_owners(o: Obj, owners: List[Obj]) -> Obj:
    """expectedShardSecret / expectedShardConfigMap / provisionOwnedControllerResources (:198-228, :754-775)."""
    return Obj(o.kind, o.namespace, o.name, o.uid, o.value, [("NexusAlgorithmTemplate", t.uid) for t in owners],
               list(o.secret_names), list(o.configmap_names))


def with_uid(o: Obj, uid: str) -> Obj:
    """expectedShardTemplate(template, uid) (:190-196): a copy with another UID (labels are not compared)."""
    return Obj(o.kind, o.namespace, o.name, uid, o.value, list(o.owner_refs), list(o.secret_names), list(o.configmap_names))
