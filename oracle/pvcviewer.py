"""Oracle (test infrastructure only): what one PVCViewer reconcile writes, and the viewer's
defaulting / validating webhook.

Follows components/pvcviewer-controller:
  controllers/pvcviewer_controller.go — Reconcile :96-146 (commonLabels :121-125),
    reconcileDeployment :149-208, reconcileService :211-250, reconcileVirtualService :252-336,
    reconcileStatus :339-369, the Affinity generateAffinity ends with :445-464;
  api/v1alpha1/pvcviewer_webhook.go — Default :70-146, validate :153-178.
Each reconcile* function Gets the object, overwrites the fields it owns and Creates (NotFound) or
Updates what it then holds: `found` below is that Get's result (None = NotFound) and the return
value is the body of the write.  The PVC Get and the Pod List behind the affinity are I/O: the
node name they lead to is a parameter.  Pinned by the assertions of
controllers/pvcviewer_controller_test.go (tests/golden/pvcviewer.json, transcribed).
"""
from .gojson import deepcopy, meta

PREFIX = "pvcviewer-"                                                                   # :49
DEFAULT_GATEWAY = "kubeflow/kubeflow-gateway"                                           # :58


class AlreadyOwned(Exception):
    """ctrl.SetControllerReference: the object has another controller."""


def common_labels(viewer):
    name = meta(viewer).get("name", "")
    return {"app.kubernetes.io/name": name, "app.kubernetes.io/instance": PREFIX + name,
            "app.kubernetes.io/part-of": "pvc-viewer"}                                  # :121-125


def owner_ref(viewer):
    m = meta(viewer)
    return {"apiVersion": "kubeflow.org/v1alpha1", "kind": "PVCViewer", "name": m.get("name", ""),
            "uid": m.get("uid", ""), "controller": True, "blockOwnerDeletion": True}


def set_controller_reference(viewer, obj):
    """controllerutil.SetControllerReference: refuse another controller, replace an entry for
    the same owner (group/kind/name), else append."""
    ref = owner_ref(viewer)
    refs = list(meta(obj).get("ownerReferences") or [])
    for r in refs:
        if r.get("controller") and r.get("uid", "") != ref["uid"]:
            raise AlreadyOwned(r.get("name", ""))
    for i, r in enumerate(refs):
        same_group = (r.get("apiVersion", "").split("/")[0] == "kubeflow.org")
        if same_group and r.get("kind") == ref["kind"] and r.get("name") == ref["name"]:
            refs[i] = ref
            break
    else:
        refs.append(ref)
    obj.setdefault("metadata", {})["ownerReferences"] = refs


def networking(viewer):
    return (viewer.get("spec") or {}).get("networking") or {}


def networking_is_zero(viewer):
    """viewer.Spec.Networking == (Networking{}) :212,:253: every field at its zero value."""
    n = networking(viewer)
    tp = n.get("targetPort")
    return (tp in (None, 0, "")) and not n.get("basePrefix") and not n.get("rewrite") and not n.get("timeout")


def node_affinity(nodename):
    return {"nodeAffinity": {"preferredDuringSchedulingIgnoredDuringExecution": [
        {"weight": 100, "preference": {"matchExpressions": [
            {"key": "kubernetes.io/hostname", "operator": "In", "values": [nodename]}]}}]}}   # :445-464


def _base_meta(viewer):
    m = {"name": PREFIX + meta(viewer).get("name", ""), "labels": common_labels(viewer)}
    if meta(viewer).get("namespace", ""):
        m["namespace"] = meta(viewer)["namespace"]
    return m


def reconcile_deployment(viewer, found=None, nodename=""):
    """reconcileDeployment :149-208: the body of the Create (found None) or Update."""
    labels = common_labels(viewer)
    spec_in = viewer.get("spec") or {}
    dep = deepcopy(found) if found is not None else {"apiVersion": "apps/v1", "kind": "Deployment",
                                                       "metadata": _base_meta(viewer)}
    old_pod = ((((dep.get("spec") or {}).get("template") or {}).get("spec")) or {})
    affinity = old_pod.get("affinity")                                                  # :168
    if spec_in.get("rwoScheduling") and found is None and nodename:                     # :170-180
        affinity = node_affinity(nodename)
    if not isinstance(dep.get("spec"), dict):
        dep["spec"] = {}
    spec = dep["spec"]
    spec["selector"] = {"matchLabels": labels}                                          # :182-184
    pod = deepcopy(spec_in.get("podSpec") or {})
    if affinity is None:
        pod.pop("affinity", None)                                                       # :196 nil pointer, omitempty
    else:
        pod["affinity"] = deepcopy(affinity)
    spec["template"] = {"metadata": {"labels": labels}, "spec": pod}                    # :185-190
    spec["strategy"] = {"type": "Recreate"}                                             # :193-195
    set_controller_reference(viewer, dep)                                               # :198
    return dep


def reconcile_service(viewer, found=None):
    """reconcileService :211-250; None when spec.networking is the zero value."""
    if networking_is_zero(viewer):
        return None
    svc = deepcopy(found) if found is not None else {"apiVersion": "v1", "kind": "Service", "metadata": _base_meta(viewer)}
    spec = svc.get("spec") if isinstance(svc.get("spec"), dict) else None
    if spec is None:
        spec = svc["spec"] = {}
    spec["type"] = "ClusterIP"
    spec["selector"] = common_labels(viewer)
    tp = networking(viewer).get("targetPort")
    spec["ports"] = [{"name": "http", "port": 80, "targetPort": 0 if tp is None else tp}]   # :232-238
    set_controller_reference(viewer, svc)
    return svc


def reconcile_virtualservice(viewer, env, found=None):
    """reconcileVirtualService :252-336; None when spec.networking is the zero value."""
    if networking_is_zero(viewer):
        return None
    name, ns = meta(viewer).get("name", ""), meta(viewer).get("namespace", "")
    n = networking(viewer)
    if found is not None:
        vs = deepcopy(found)
    else:
        vs = {"apiVersion": "networking.istio.io/v1alpha3", "kind": "VirtualService",
              "metadata": {"name": PREFIX + name, "namespace": ns, "labels": common_labels(viewer)}}   # :257-267
    prefix = "%s/%s/%s/" % (n.get("basePrefix") or "", ns, name)                        # :277
    rewrite = n.get("rewrite") or prefix                                                # :278-281
    gateway = env.get("ISTIO_GATEWAY") or DEFAULT_GATEWAY                               # :289-292
    vs["spec"] = {"hosts": ["*"], "gateways": [gateway],
                  "http": [{"match": [{"uri": {"prefix": prefix}}], "rewrite": {"uri": rewrite},
                            "route": [{"destination": {"host": "%s%s.%s.svc.cluster.local" % (PREFIX, name, ns),
                                                       "port": {"number": 80}}}],
                            "timeout": n.get("timeout") or None}]}                      # :294-322 (nil *string -> null)
    set_controller_reference(viewer, vs)
    return vs


def viewer_objects(viewer, env, found=(None, None, None), nodename=""):
    """(Deployment, Service or None, VirtualService or None) one Reconcile writes."""
    return (reconcile_deployment(viewer, found[0], nodename), reconcile_service(viewer, found[1]),
            reconcile_virtualservice(viewer, env, found[2]))


def reconcile_status(viewer, deployment):
    """reconcileStatus :339-369: the viewer's new status; deployment None = the Get failed."""
    status = deepcopy(viewer.get("status") or {})
    if not networking_is_zero(viewer):
        status["url"] = "%s/%s/%s/" % (networking(viewer).get("basePrefix") or "", meta(viewer).get("namespace", ""),
                                       meta(viewer).get("name", ""))
    else:
        status.pop("url", None)
    if deployment is None:
        status["ready"] = False
        if not status.get("conditions"):
            status.pop("conditions", None)
        return status
    replicas = (deployment.get("spec") or {}).get("replicas")
    if replicas is None:
        raise ValueError("nil pointer dereference: Spec.Replicas")                      # :357
    status["ready"] = replicas == (deployment.get("status") or {}).get("readyReplicas", 0)
    dconds = (deployment.get("status") or {}).get("conditions") or []
    if dconds:                                                                          # :359-364
        conds = list(status.get("conditions") or [])
        if not conds or conds[-1] != dconds[0]:
            conds.append(deepcopy(dconds[0]))
        status["conditions"] = conds
    if not status.get("conditions"):
        status.pop("conditions", None)                                                  # omitempty
    return status


# ---- webhook -----------------------------------------------------------------------------------
def pod_spec_is_zero(ps):
    """reflect.DeepEqual(r.Spec.PodSpec, corev1.PodSpec{}) on the decoded struct: absent, null,
    {} or only nulls.  (An empty list is a non-nil slice and makes it non-zero.)"""
    return ps is None or (isinstance(ps, dict) and all(v is None for v in ps.values()))


def default_container(viewer):
    n = networking(viewer)
    base = "%s/%s/%s/" % (n.get("basePrefix") or "", meta(viewer).get("namespace", ""), meta(viewer).get("name", ""))
    return {"name": "pvcviewer", "image": "filebrowser/filebrowser:latest",
            "ports": [{"containerPort": 8080, "protocol": "TCP"}],
            "env": [{"name": "FB_ADDRESS", "value": "0.0.0.0"}, {"name": "FB_PORT", "value": "8080"},
                    {"name": "FB_DATABASE", "value": "/tmp/filebrowser.db"}, {"name": "FB_NOAUTH", "value": "true"},
                    {"name": "FB_BASEURL", "value": base}],
            "workingDir": "/data", "volumeMounts": [{"name": "viewer-volume", "mountPath": "/data"}]}   # :95-130


def default(viewer, default_pod_spec=None):
    """Default :70-146.  default_pod_spec: the PodSpec of the defaults file (DEFAULT_POD_SPEC_PATH),
    None when the variable is not set."""
    out = deepcopy(viewer)
    if not isinstance(out.get("spec"), dict):
        out["spec"] = {}
    spec = out["spec"]
    if not pod_spec_is_zero(spec.get("podSpec")):
        return out
    if default_pod_spec is not None and not pod_spec_is_zero(default_pod_spec):         # :88-91
        ps = deepcopy(default_pod_spec)
    else:
        ps = {"containers": [default_container(viewer)]}
    ps["volumes"] = list(ps.get("volumes") or []) + [
        {"name": "viewer-volume", "persistentVolumeClaim": {"claimName": spec.get("pvc", "") or ""}}]   # :135-143
    spec["podSpec"] = ps
    return out


def validate(viewer):
    """validate :153-178: None or the error text."""
    spec = viewer.get("spec") or {}
    pvc = spec.get("pvc") or ""
    if len(pvc) == 0:
        return "PVC name must be specified"
    if pod_spec_is_zero(spec.get("podSpec")):
        return "PodSpec must be specified"
    for v in (spec.get("podSpec") or {}).get("volumes") or []:
        c = v.get("persistentVolumeClaim")
        if isinstance(c, dict) and c.get("claimName", "") == pvc:
            return None
    return "PVC %s must be used in the podSpec" % pvc
