"""Oracle: profile-controller desired objects (test infrastructure only).

Follows components/profile-controller/controllers/profile_controller.go:
Reconcile :105-334 (the objects it builds at :127-137, :211-246, :253-262),
getAuthorizationPolicy :419-504, updateIstioAuthorizationPolicy :509-520,
updateServiceAccount :592-637, setNamespaceLabels :754-772.  Golden:
profile_controller_test.go:25-111 (setNamespaceLabels, 3 cases).  The JSON shape
of the Istio AuthorizationPolicy spec (proto JSON, ALLOW omitted as the zero
enum) is recalled from istio.io/api, not vendored: parity unpinned.
"""
from .gojson import deepcopy, meta

DEFAULTS = {
    "USERID_HEADER": "x-goog-authenticated-user-email",          # main.go:75
    "USERID_PREFIX": "accounts.google.com:",                     # main.go:76
    "NOTEBOOK_CONTROLLER_PRINCIPAL": "cluster.local/ns/kubeflow/sa/notebook-controller-service-account",
    "ISTIO_INGRESS_GATEWAY_PRINCIPAL": "cluster.local/ns/istio-system/sa/istio-ingressgateway-service-account",
    "KFP_UI_PRINCIPAL": "cluster.local/ns/kubeflow/sa/ml-pipeline-ui",
}
PROFILE_SLOTS = ("namespace", "authorizationpolicy", "sa_editor", "rb_editor", "sa_viewer", "rb_viewer",
                 "rb_admin", "resourcequota")


def set_namespace_labels(labels, new_labels):
    """setNamespaceLabels, :754-772, on a labels dict (None = nil map).  Returns the dict."""
    if labels is None:
        labels = {}
    for k, v in new_labels.items():
        if len(v) == 0:
            labels.pop(k, None)
        elif k not in labels:
            labels[k] = v
    return labels


def _env(env, k):
    return env.get(k) or DEFAULTS[k]       # GetEnvDefault: unset or empty -> default


def _owner_ref(profile):
    m = meta(profile)
    return {"apiVersion": "kubeflow.org/v1", "kind": "Profile", "name": m.get("name", ""),
            "uid": m.get("uid", ""), "controller": True, "blockOwnerDeletion": True}


def profile_objects(profile, env, default_labels):
    """The 8 desired objects of one Profile reconcile, PROFILE_SLOTS order; the
    ResourceQuota slot is None when spec.resourceQuotaSpec.hard is empty (:253)."""
    name = meta(profile).get("name", "")
    spec = profile.get("spec") or {}
    owner = spec.get("owner") or {}
    owner_name = owner.get("name", "")
    ref = _owner_ref(profile)

    def md(n, namespaced=True, annotations=None, labels=None):
        m = {"name": n}
        if namespaced:
            m["namespace"] = name
        if annotations:
            m["annotations"] = annotations
        if labels:
            m["labels"] = labels
        m["ownerReferences"] = [deepcopy(ref)]
        return m

    ns = {"apiVersion": "v1", "kind": "Namespace",
          "metadata": md(name, False, {"owner": owner_name},
                         set_namespace_labels({"istio-injection": "enabled"}, default_labels))}
    authz = {"apiVersion": "security.istio.io/v1beta1", "kind": "AuthorizationPolicy",
             "metadata": md("ns-owner-access-istio", True, {"user": owner_name, "role": "admin"}),
             "spec": {"rules": [
                 {"from": [{"source": {"principals": [_env(env, "ISTIO_INGRESS_GATEWAY_PRINCIPAL"),
                                                      _env(env, "KFP_UI_PRINCIPAL")]}}],
                  "when": [{"key": "request.headers[%s]" % _env(env, "USERID_HEADER"),
                            "values": [_env(env, "USERID_PREFIX") + owner_name]}]},
                 {"when": [{"key": "source.namespace", "values": [name]}]},
                 {"to": [{"operation": {"paths": ["/healthz", "/metrics", "/wait-for-drain"]}}]},
                 {"from": [{"source": {"principals": [_env(env, "NOTEBOOK_CONTROLLER_PRINCIPAL")]}}],
                  "to": [{"operation": {"methods": ["GET"], "paths": ["*/api/kernels"]}}]}]}}

    def sa(n):
        return {"apiVersion": "v1", "kind": "ServiceAccount", "metadata": md(n)}

    def rb(n, role, subjects, annotations=None):
        return {"apiVersion": "rbac.authorization.k8s.io/v1", "kind": "RoleBinding",
                "metadata": md(n, True, annotations),
                "roleRef": {"apiGroup": "rbac.authorization.k8s.io", "kind": "ClusterRole", "name": role},
                "subjects": subjects}

    def sa_subject(n):
        return [{"kind": "ServiceAccount", "name": n, "namespace": name}]

    quota = None
    rq = spec.get("resourceQuotaSpec") or {}
    if rq.get("hard"):
        quota = {"apiVersion": "v1", "kind": "ResourceQuota", "metadata": md("kf-resource-quota"),
                 "spec": deepcopy(rq)}
    return [ns, authz, sa("default-editor"), rb("default-editor", "kubeflow-edit", sa_subject("default-editor")),
            sa("default-viewer"), rb("default-viewer", "kubeflow-view", sa_subject("default-viewer")),
            rb("namespaceAdmin", "kubeflow-admin", [deepcopy(owner)], {"user": owner_name, "role": "admin"}),
            quota]


def kfam_binding_name(user_kind, user_name, roleref_kind, roleref_name):
    """getBindingName, components/access-management/kfam/bindings.go:61-78: every run of
    characters outside [a-zA-Z0-9] becomes one '-', the four parts are joined by '-' and
    lower-cased (url.QueryEscape is the identity on what the first replacement leaves)."""
    import re
    reg = re.compile(r"[^a-zA-Z0-9]+")
    raw = "-".join([user_kind, reg.sub("-", user_name), roleref_kind, roleref_name]).lower()
    return reg.sub("-", raw)
