"""Oracle: odh-notebook-controller generators and webhook mutations (test
infrastructure only).

Follows components/odh-notebook-controller/controllers/{notebook_controller,
notebook_network,notebook_oauth,notebook_route,notebook_rbac,notebook_webhook,
notebook_webhook_utils}.go on JSON trees.  Golden vectors:
notebook_controller_test.go:61-87 (Route), :436-477 (ctrl NetworkPolicy),
:602-658 + :1002-1086 (mutated OAuth Notebook), :728-747 (TLS Service),
:798-823 (OAuth Route), :986-1000 (ServiceAccount), :1088-1119 (OAuth
NetworkPolicy), :142-149 (ownerReference).
"""
from .gojson import deepcopy, get_map, has_annotation, meta, parse_bool
from .nbc import STOP_ANNOTATION, json_equal

ANNOTATION_INJECT_OAUTH = "notebooks.opendatahub.io/inject-oauth"          # notebook_controller.go:51
ANNOTATION_SERVICE_MESH = "opendatahub.io/service-mesh"                    # :52
ANNOTATION_VALUE_RECONCILIATION_LOCK = "odh-notebook-controller-lock"      # :53
ANNOTATION_LOGOUT_URL = "notebooks.opendatahub.io/oauth-logout-url"        # :54
ANNOTATION_UPDATE_PENDING = "notebooks.opendatahub.io/update-pending"      # notebook_webhook.go:288
ANNOTATION_NOTEBOOK_RESTART = "notebooks.opendatahub.io/notebook-restart"
NOTEBOOK_PORT = 8888                                                       # notebook_network.go:34-37
NOTEBOOK_OAUTH_PORT = 8443
OAUTH_SERVICE_PORT = 443                                                   # notebook_oauth.go:36-37
OAUTH_SERVICE_PORT_NAME = "oauth-proxy"
OAUTH_PROXY_IMAGE = ("registry.redhat.io/openshift4/ose-oauth-proxy@sha256:"
                     "4f8d66597feeb32bb18699326029f9a71a5aca4a57679d636b876377c2e95695")  # :41


def _name_ns(nb):
    m = meta(nb)
    return m.get("name", ""), m.get("namespace", "")


def _ometa(name, ns, labels=None, annotations=None):
    m = {"name": name}
    if ns:
        m["namespace"] = ns
    if labels:
        m["labels"] = labels
    if annotations:
        m["annotations"] = annotations
    return m


def oauth_injection_is_enabled(m):
    """OAuthInjectionIsEnabled, notebook_controller.go:86-93."""
    v = get_map(m, "annotations").get(ANNOTATION_INJECT_OAUTH, "")
    return parse_bool(v)[0] if v != "" else False


def service_mesh_is_enabled(m):
    """ServiceMeshIsEnabled, notebook_controller.go:97-104."""
    v = get_map(m, "annotations").get(ANNOTATION_SERVICE_MESH, "")
    return parse_bool(v)[0] if v != "" else False


def reconciliation_lock_is_enabled(m):
    """ReconciliationLockIsEnabled, notebook_controller.go:108-114."""
    v = get_map(m, "annotations").get(STOP_ANNOTATION, "")
    return v == ANNOTATION_VALUE_RECONCILIATION_LOCK if v != "" else False


def compare_notebooks(a, b):
    """CompareNotebooks, notebook_controller.go:78-82."""
    return (json_equal(meta(a).get("labels"), meta(b).get("labels")) and
            json_equal(meta(a).get("annotations"), meta(b).get("annotations")) and
            json_equal(a.get("spec"), b.get("spec")))


# ---------------------------------------------------------------- generators


def new_notebook_network_policy(nb, ctrl_namespace):
    """NewNotebookNetworkPolicy, notebook_network.go:131-173."""
    name, ns = _name_ns(nb)
    return {"apiVersion": "networking.k8s.io/v1", "kind": "NetworkPolicy",
            "metadata": _ometa(name + "-ctrl-np", ns),
            "spec": {"podSelector": {"matchLabels": {"notebook-name": name}},
                     "ingress": [{"ports": [{"protocol": "TCP", "port": NOTEBOOK_PORT}],
                                  "from": [{"namespaceSelector": {"matchLabels": {
                                      "kubernetes.io/metadata.name": ctrl_namespace}}}]}],
                     "policyTypes": ["Ingress"]}}


def new_oauth_network_policy(nb):
    """NewOAuthNetworkPolicy, notebook_network.go:176-210."""
    name, ns = _name_ns(nb)
    return {"apiVersion": "networking.k8s.io/v1", "kind": "NetworkPolicy",
            "metadata": _ometa(name + "-oauth-np", ns),
            "spec": {"podSelector": {"matchLabels": {"notebook-name": name}},
                     "ingress": [{"ports": [{"protocol": "TCP", "port": NOTEBOOK_OAUTH_PORT}]}],
                     "policyTypes": ["Ingress"]}}


def compare_notebook_network_policies(a, b):
    """CompareNotebookNetworkPolicies, notebook_network.go:124-128."""
    return json_equal(meta(a).get("labels"), meta(b).get("labels")) and \
        json_equal(a.get("spec"), b.get("spec"))


def new_notebook_route(nb):
    """NewNotebookRoute, notebook_route.go:34-62."""
    name, ns = _name_ns(nb)
    return {"apiVersion": "route.openshift.io/v1", "kind": "Route",
            "metadata": _ometa(name, ns, {"notebook-name": name}),
            "spec": {"to": {"kind": "Service", "name": name, "weight": 100},
                     "port": {"targetPort": "http-" + name},
                     "tls": {"termination": "edge", "insecureEdgeTerminationPolicy": "Redirect"},
                     "wildcardPolicy": "None"},
            "status": {"ingress": []}}


def new_notebook_oauth_route(nb):
    """NewNotebookOAuthRoute, notebook_oauth.go:253-259."""
    name, _ = _name_ns(nb)
    r = new_notebook_route(nb)
    r["spec"]["to"]["name"] = name + "-tls"
    r["spec"]["port"]["targetPort"] = OAUTH_SERVICE_PORT_NAME
    r["spec"]["tls"]["termination"] = "reencrypt"
    return r


def compare_notebook_routes(a, b):
    """CompareNotebookRoutes, notebook_route.go:65-72 (host ignored)."""
    sa, sb = deepcopy(a.get("spec") or {}), deepcopy(b.get("spec") or {})
    sa.pop("host", None)
    sb.pop("host", None)
    return json_equal(meta(a).get("labels"), meta(b).get("labels")) and json_equal(sa, sb)


def new_notebook_service_account(nb):
    """NewNotebookServiceAccount, notebook_oauth.go:49-64."""
    name, ns = _name_ns(nb)
    return {"apiVersion": "v1", "kind": "ServiceAccount",
            "metadata": _ometa(name, ns, {"notebook-name": name}, {
                "serviceaccounts.openshift.io/oauth-redirectreference.first":
                    '{"kind":"OAuthRedirectReference","apiVersion":"v1",'
                    '"reference":{"kind":"Route","name":"' + name + '"}}'})}


def new_notebook_oauth_service(nb):
    """NewNotebookOAuthService, notebook_oauth.go:116-140."""
    name, ns = _name_ns(nb)
    return {"apiVersion": "v1", "kind": "Service",
            "metadata": _ometa(name + "-tls", ns, {"notebook-name": name},
                               {"service.beta.openshift.io/serving-cert-secret-name": name + "-tls"}),
            "spec": {"ports": [{"name": OAUTH_SERVICE_PORT_NAME, "protocol": "TCP",
                                "port": OAUTH_SERVICE_PORT, "targetPort": OAUTH_SERVICE_PORT_NAME}],
                     "selector": {"statefulset": name}}}


def new_notebook_oauth_secret(nb, cookie_secret):
    """NewNotebookOAuthSecret, notebook_oauth.go:190-210.  The reference draws
    16 random bytes (crypto/rand) and stores base64(base64(seed)); the only
    pinned property is len == 32 (notebook_controller_test.go:782).  The
    caller supplies the cookie so the function stays deterministic."""
    name, ns = _name_ns(nb)
    return {"apiVersion": "v1", "kind": "Secret",
            "metadata": _ometa(name + "-oauth-config", ns, {"notebook-name": name}),
            "stringData": {"cookie_secret": cookie_secret}}


def cookie_secret_from_seed(seed16):
    import base64
    assert len(seed16) == 16
    return base64.b64encode(base64.b64encode(seed16)).decode()


def new_role_binding(nb, rolebinding_name, role_ref_kind, role_ref_name):
    """NewRoleBinding, notebook_rbac.go:36-58."""
    name, ns = _name_ns(nb)
    return {"apiVersion": "rbac.authorization.k8s.io/v1", "kind": "RoleBinding",
            "metadata": _ometa(rolebinding_name, ns, {"notebook-name": name}),
            "subjects": [dict({"kind": "ServiceAccount", "name": name}, **({"namespace": ns} if ns else {}))],
            "roleRef": {"apiGroup": "rbac.authorization.k8s.io", "kind": role_ref_kind,
                        "name": role_ref_name}}


def elyra_role_binding(nb):
    """ReconcileRoleBindings, notebook_rbac.go:144-153."""
    name, _ = _name_ns(nb)
    return new_role_binding(nb, "elyra-pipelines-" + name, "Role", "ds-pipeline-user-access-dspa")


ODH_SLOTS = ("ctrl_np", "oauth_np", "route", "service_account", "service", "secret", "rolebinding")


def reconcile_objects(nb, ctrl_namespace, rbac=False, cookie_secret="", owner_ref=True):
    """The desired objects OpenshiftNotebookReconciler.Reconcile builds for one
    Notebook (notebook_controller.go:188-232), one slot per ODH_SLOTS entry,
    None where the reference builds nothing:
      ctrl NetworkPolicy always (:46 of notebook_network.go); OAuth NetworkPolicy
      unless service mesh (:54-57); RoleBinding when SET_PIPELINE_RBAC is "true"
      and the Role exists (`rbac`); without service mesh, OAuth on -> SA, Service,
      Secret, OAuth Route, OAuth off -> plain Route.
    Each object carries the controller ownerReference (apiVersion kubeflow.org/v1)
    the Reconcile* helpers add before Create."""
    from .nbc import set_controller_reference
    m = meta(nb)
    mesh = service_mesh_is_enabled(m)
    oauth = oauth_injection_is_enabled(m)
    out = [new_notebook_network_policy(nb, ctrl_namespace),
           None if mesh else new_oauth_network_policy(nb),
           None, None, None, None,
           elyra_role_binding(nb) if rbac else None]
    if not mesh:
        if oauth:
            out[2] = new_notebook_oauth_route(nb)
            out[3] = new_notebook_service_account(nb)
            out[4] = new_notebook_oauth_service(nb)
            out[5] = new_notebook_oauth_secret(nb, cookie_secret)
        else:
            out[2] = new_notebook_route(nb)
    if owner_ref:
        for o in out:
            if o is not None:
                set_controller_reference(nb, o, "kubeflow.org/v1")
    return out


def write_decision(desired, found):
    """What the ODH reconciler of one object writes, given what it Got:
    (None, None) | ("create", obj) | ("update", obj).
      Route           reconcileRoute, notebook_route.go:76-139 (CompareNotebookRoutes :65-72)
      NetworkPolicy   reconcileNetworkPolicy, notebook_network.go:66-122 (:124-128)
      ServiceAccount / Service / Secret   notebook_oauth.go:77-113, :151-187, :214-250:
                      Create when NotFound, otherwise left alone
      RoleBinding     reconcileRoleBinding, notebook_rbac.go:89-140: Update(desired) when the
                      subjects differ; the ownerReference is only set on the Create branch.
    `desired` carries the ownerReference (reconcile_objects), `found` is None for NotFound."""
    if desired is None:
        return None, None
    if found is None:
        return "create", desired
    kind = desired.get("kind")
    if kind in ("Route", "NetworkPolicy"):
        same = compare_notebook_routes(desired, found) if kind == "Route" \
            else compare_notebook_network_policies(desired, found)
        if same:
            return None, None
        out = deepcopy(found)
        if desired.get("spec") is None:
            out.pop("spec", None)
        else:
            out["spec"] = deepcopy(desired["spec"])
        md = out.setdefault("metadata", {})
        if meta(desired).get("labels") is None:
            md.pop("labels", None)
        else:
            md["labels"] = deepcopy(meta(desired)["labels"])
        return "update", out
    if kind == "RoleBinding":
        if json_equal(desired.get("subjects"), found.get("subjects")):
            return None, None
        out = deepcopy(desired)
        out.get("metadata", {}).pop("ownerReferences", None)
        return "update", out
    return None, None


# ---------------------------------------------------------------- webhook mutations


def inject_reconciliation_lock(m):
    """InjectReconciliationLock, notebook_webhook.go:63-72.  Mutates metadata."""
    if m.get("annotations") is not None:
        m["annotations"][STOP_ANNOTATION] = ANNOTATION_VALUE_RECONCILIATION_LOCK
    else:
        m["annotations"] = {STOP_ANNOTATION: ANNOTATION_VALUE_RECONCILIATION_LOCK}


def oauth_proxy_container(name, image, logout_url=""):
    """The sidecar of InjectOAuthProxy, notebook_webhook.go:78-165."""
    probe = {"httpGet": {"path": "/oauth/healthz", "port": OAUTH_SERVICE_PORT_NAME, "scheme": "HTTPS"}}
    args = [
        "--provider=openshift", "--https-address=:8443", "--http-address=",
        "--openshift-service-account=" + name,
        "--cookie-secret-file=/etc/oauth/config/cookie_secret", "--cookie-expire=24h0m0s",
        "--tls-cert=/etc/tls/private/tls.crt", "--tls-key=/etc/tls/private/tls.key",
        "--upstream=http://localhost:8888",
        "--upstream-ca=/var/run/secrets/kubernetes.io/serviceaccount/ca.crt",
        "--email-domain=*", "--skip-provider-button",
        '--openshift-sar={"verb":"get","resource":"notebooks","resourceAPIGroup":"kubeflow.org",'
        '"resourceName":"' + name + '","namespace":"$(NAMESPACE)"}',
    ]
    if logout_url != "":
        args.append("--logout-url=" + logout_url)
    res = {"cpu": "100m", "memory": "64Mi"}
    return {
        "name": "oauth-proxy", "image": image, "args": args,
        "ports": [{"name": OAUTH_SERVICE_PORT_NAME, "containerPort": 8443, "protocol": "TCP"}],
        "env": [{"name": "NAMESPACE", "valueFrom": {"fieldRef": {"fieldPath": "metadata.namespace"}}}],
        "resources": {"limits": dict(res), "requests": dict(res)},
        "volumeMounts": [{"name": "oauth-config", "mountPath": "/etc/oauth/config"},
                         {"name": "tls-certificates", "mountPath": "/etc/tls/private"}],
        "livenessProbe": dict(probe, initialDelaySeconds=30, timeoutSeconds=1, periodSeconds=5,
                              successThreshold=1, failureThreshold=3),
        "readinessProbe": dict(probe, initialDelaySeconds=5, timeoutSeconds=1, periodSeconds=5,
                               successThreshold=1, failureThreshold=3),
        "imagePullPolicy": "Always",
    }


def _replace_or_append(lst, item):
    for i, x in enumerate(lst):
        if (x or {}).get("name", "") == item["name"]:
            lst[i] = item
            return
    lst.append(item)


def inject_oauth_proxy(nb, image=OAUTH_PROXY_IMAGE):
    """InjectOAuthProxy, notebook_webhook.go:76-231.  Mutates nb."""
    name, _ = _name_ns(nb)
    logout = get_map(meta(nb), "annotations").get(ANNOTATION_LOGOUT_URL, "")
    spec = nb.setdefault("spec", {}).setdefault("template", {}).setdefault("spec", {})
    if spec.get("containers") is None:
        spec["containers"] = []
    _replace_or_append(spec["containers"], oauth_proxy_container(name, image, logout))   # :168-179
    if spec.get("volumes") is None:
        spec["volumes"] = []
    _replace_or_append(spec["volumes"], {"name": "oauth-config", "secret": {
        "secretName": name + "-oauth-config", "defaultMode": 420}})                      # :183-203
    _replace_or_append(spec["volumes"], {"name": "tls-certificates", "secret": {
        "secretName": name + "-tls", "defaultMode": 420}})                               # :207-226
    spec["serviceAccountName"] = name                                                    # :229


CERT_MOUNT_PATH = "/etc/pki/tls/custom-certs/ca-bundle.crt"
CERT_ENV_VARS = ["PIP_CERT", "REQUESTS_CA_BUNDLE", "SSL_CERT_FILE", "PIPELINES_SSL_SA_CERTS",
                 "GIT_SSL_CAINFO"]   # Go iterates a map here (:476): order is random in the reference


def inject_cert_config(nb, config_map_name="workbench-trusted-ca-bundle"):
    """InjectCertConfig, notebook_webhook.go:421-535.  Mutates nb.

    The reference appends the missing env vars in Go map order (random); this
    restatement appends them in the order of CERT_ENV_VARS and parity checks
    compare the env list as a set for those five names."""
    name, _ = _name_ns(nb)
    spec = nb.setdefault("spec", {}).setdefault("template", {}).setdefault("spec", {})
    if spec.get("volumes") is None:
        spec["volumes"] = []
    _replace_or_append(spec["volumes"], {"name": "trusted-ca", "configMap": {
        "name": config_map_name, "items": [{"key": "ca-bundle.crt", "path": "ca-bundle.crt"}],
        "optional": True}})                                                              # :440-467
    for c in spec.get("containers") or []:
        if c.get("name", "") != name:                                                    # :472
            continue
        have = {e.get("name", "") for e in c.get("env") or []}
        new = [{"name": k, "value": CERT_MOUNT_PATH} for k in CERT_ENV_VARS if k not in have]
        if new:                                                                          # :494-496
            c["env"] = (c.get("env") or []) + new
        # an existing var with another value is NOT updated: the Go loop writes to a copy (:482-484)
        mounts = c.get("volumeMounts")
        if mounts is None:
            mounts = []
        _replace_or_append(mounts, {"name": "trusted-ca", "readOnly": True,
                                    "mountPath": CERT_MOUNT_PATH, "subPath": "ca-bundle.crt"})  # :498-517
        c["volumeMounts"] = mounts
        break


def first_difference(a, b, path="{v1.PodSpec}"):
    """getStructDiff / FirstDifferenceReporter, notebook_webhook_utils.go:27-80,
    restated on JSON trees.  go-cmp walks struct fields in declaration order and
    renders Go field names; this walks JSON keys in sorted order and renders JSON
    names, so only the "<path>: <x> != <y>" shape is kept
    (notebook_webhook_utils_test.go:36,55 pin that shape)."""
    if type(a) is not type(b):
        return "%s: %s != %s" % (path, a, b)
    if isinstance(a, dict):
        for k in sorted(set(a) | set(b)):
            if k not in a or k not in b:
                return "%s.%s: %s != %s" % (path, k, a.get(k), b.get(k))
            d = first_difference(a[k], b[k], path + "." + k)
            if d:
                return d
        return ""
    if isinstance(a, list):
        if len(a) != len(b):
            return "%s: len %d != %d" % (path, len(a), len(b))
        for i, (x, y) in enumerate(zip(a, b)):
            d = first_difference(x, y, "%s[%d]" % (path, i))
            if d:
                return d
        return ""
    return "" if a == b else "%s: %s != %s" % (path, a, b)


def semantic_equal(a, b):
    """equality.Semantic.DeepEqual on JSON trees: nil and empty slices/maps are
    equal; quantities are compared as text here."""
    def norm(x):
        if isinstance(x, dict):
            return {k: norm(v) for k, v in x.items() if v not in (None, [], {})}
        if isinstance(x, list):
            return [norm(v) for v in x]
        return x
    return json_equal(norm(a), norm(b))


class Denied(Exception):
    """admission.Denied (notebook_webhook.go:274)."""


def mutate_notebook(nb, operation, old=None, image=OAUTH_PROXY_IMAGE, ca_bundle=False):
    """NotebookWebhook.Handle, notebook_webhook.go:234-302, minus the client
    reads: `ca_bundle` says whether CheckAndMountCACertBundle found its
    ConfigMaps (:373-419); SetContainerImageFromRegistry (:541-640) needs a
    live ImageStream list and is out of the pure path.

    Returns (mutated notebook tree, update_pending_reason or "")."""
    updated = deepcopy(nb)
    nb = deepcopy(nb)
    m = nb.setdefault("metadata", {})
    if operation == "CREATE":
        inject_reconciliation_lock(m)                                                    # :248-254
    if operation in ("CREATE", "UPDATE") and ca_bundle:
        inject_cert_config(nb)                                                           # :265
    if oauth_injection_is_enabled(m):                                                    # :272-280
        if service_mesh_is_enabled(m):
            raise Denied("Cannot have both %s and %s set to true. Pick one." %
                         (ANNOTATION_SERVICE_MESH, ANNOTATION_INJECT_OAUTH))
        inject_oauth_proxy(nb, image)
    reason = ""
    # maybeRestartRunningNotebook, :314-370
    if not (operation == "CREATE" or has_annotation(m, STOP_ANNOTATION) or
            has_annotation(m, ANNOTATION_NOTEBOOK_RESTART)):
        def pspec(x):
            return (((x or {}).get("spec") or {}).get("template") or {}).get("spec") or {}
        if semantic_equal(pspec(old), pspec(updated)) and \
                not semantic_equal(pspec(old), pspec(nb)):                               # :353-362
            reason = first_difference(pspec(nb), pspec(updated))                         # :366
            nb["spec"]["template"]["spec"] = deepcopy(pspec(updated))                    # :368
    if reason:                                                                           # :289-293
        m.setdefault("annotations", {})[ANNOTATION_UPDATE_PENDING] = reason
    elif isinstance(m.get("annotations"), dict):
        m["annotations"].pop(ANNOTATION_UPDATE_PENDING, None)
    return nb, reason
