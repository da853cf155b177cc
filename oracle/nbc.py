"""Oracle: notebook-controller pure functions (test infrastructure only).

Follows components/notebook-controller/controllers/notebook_controller.go and
components/common/reconcilehelper/util.go of the reference, on JSON trees.

`env` is a dict standing for the process environment the reference reads with
os.Getenv / os.LookupEnv (USE_ISTIO, ISTIO_GATEWAY, ISTIO_HOST,
CLUSTER_DOMAIN, ADD_FSGROUP); a key that is absent means "unset".

Output convention (DESIGN.md "canonical form"): typed objects are emitted the
way Go's encoding/json would emit the fields the generator sets, with
omitempty applied, plus apiVersion/kind; zero-struct noise Go also emits
(creationTimestamp:null, status:{}) is left out; the pass-through PodSpec is
copied as it came.  Parity is compared on gojson.canon() of these trees.
generateStatefulSet/Service/VirtualService and Copy*Fields have no asserting
test in the reference: parity for them is pinned by the source text only.
"""
import json

from .gojson import deepcopy, get_map, has_annotation, meta

DEFAULT_CONTAINER_PORT = 8888          # notebook_controller.go:49
DEFAULT_SERVING_PORT = 80              # :50
ANNOTATION_REWRITE_URI = "notebooks.kubeflow.org/http-rewrite-uri"             # :51
ANNOTATION_HEADERS_REQUEST_SET = "notebooks.kubeflow.org/http-headers-request-set"  # :52
ANNOTATION_NOTEBOOK_RESTART = "notebooks.opendatahub.io/notebook-restart"      # :53
WORKBENCH_LABEL = "opendatahub.io/workbenches"                                 # :54
PREFIX_ENV_VAR = "NB_PREFIX"           # :56
DEFAULT_FS_GROUP = 100                 # :60
STOP_ANNOTATION = "kubeflow-resource-stopped"  # culling_controller.go:50


class ReferencePanic(Exception):
    """The Go code would panic here (index out of range / nil dereference)."""


def json_equal(a, b):
    """reflect.DeepEqual on JSON trees: bool is not int, [] is not null."""
    if type(a) is not type(b):
        return False
    if isinstance(a, dict):
        # a member whose value is null is a nil field after decoding: same as absent
        ka = {k for k, v in a.items() if v is not None}
        kb = {k for k, v in b.items() if v is not None}
        if ka != kb:
            return False
        return all(json_equal(a[k], b[k]) for k in ka)
    if isinstance(a, list):
        return len(a) == len(b) and all(json_equal(x, y) for x, y in zip(a, b))
    return a == b


def _name_ns(nb):
    m = meta(nb)
    return m.get("name", ""), m.get("namespace", "")


def _obj_meta(name, ns):
    m = {"name": name}
    if ns:
        m["namespace"] = ns
    return m


def _pod_spec(nb):
    return (((nb.get("spec") or {}).get("template") or {}).get("spec")) or {}


def set_controller_reference(owner, obj, api_version="kubeflow.org/v1beta1", kind="Notebook"):
    """ctrl.SetControllerReference (notebook_controller.go:141,183,628).

    Semantics recalled from controller-runtime v0.17.0;
    ownerRef shape pinned by odh .../notebook_controller_test.go:142-149."""
    m = meta(owner)
    ref = {"apiVersion": api_version, "kind": kind, "name": m.get("name", ""),
           "uid": m.get("uid", ""), "controller": True, "blockOwnerDeletion": True}
    refs = obj.setdefault("metadata", {}).setdefault("ownerReferences", [])
    for i, r in enumerate(refs):
        if r.get("controller"):
            if (r.get("apiVersion"), r.get("kind"), r.get("name")) != (api_version, kind, ref["name"]):
                raise ValueError("object is already owned by another controller")
            refs[i] = ref
            return obj
    refs.append(ref)
    return obj


def set_prefix_env_var(nb, container):
    """setPrefixEnvVar, notebook_controller.go:392-406.

    Quirk kept: when NB_PREFIX already exists the Go loop assigns to a copy of
    the range variable and returns, so the existing value stays (:395-400)."""
    name, ns = _name_ns(nb)
    prefix = "/notebook/" + ns + "/" + name
    for e in container.get("env") or []:
        if e.get("name", "") == PREFIX_ENV_VAR:
            return
    env = container.get("env")
    if env is None:
        env = []
    env.append({"name": PREFIX_ENV_VAR, "value": prefix})
    container["env"] = env


def generate_statefulset(nb, env):
    """generateStatefulSet, notebook_controller.go:408-484."""
    name, ns = _name_ns(nb)
    replicas = 0 if has_annotation(meta(nb), STOP_ANNOTATION) else 1   # :409-412

    labels = {"statefulset": name, "notebook-name": name, WORKBENCH_LABEL: "true"}  # :428-432
    for k, v in get_map(meta(nb), "labels").items():                  # :441-444
        labels[k] = v
    annotations = {}
    for k, v in get_map(meta(nb), "annotations").items():             # :447-452
        if "kubectl" not in k and "notebook" not in k:
            annotations[k] = v

    pod_spec = deepcopy(_pod_spec(nb))                                # :435
    containers = pod_spec.get("containers")
    if not containers:
        raise ReferencePanic("index out of range: Containers[0]")     # :455
    c0 = containers[0]
    if not c0.get("workingDir"):                                      # :456-458 (null decodes to "")
        c0["workingDir"] = "/home/jovyan"
    if c0.get("ports") is None:                                       # :459-467
        c0["ports"] = [{"name": "notebook-port",
                        "containerPort": DEFAULT_CONTAINER_PORT, "protocol": "TCP"}]
    set_prefix_env_var(nb, c0)                                        # :469
    if "ADD_FSGROUP" not in env or env["ADD_FSGROUP"] == "true":      # :475
        if pod_spec.get("securityContext") is None:                   # :476-481
            pod_spec["securityContext"] = {"fsGroup": DEFAULT_FS_GROUP}
    if "securityContext" in pod_spec and pod_spec["securityContext"] is None:
        del pod_spec["securityContext"]       # a nil pointer is not re-encoded (omitempty)

    tmeta = {"labels": labels}
    if annotations:                                                    # omitempty
        tmeta["annotations"] = annotations
    return {
        "apiVersion": "apps/v1", "kind": "StatefulSet",
        "metadata": _obj_meta(name, ns),
        "spec": {
            "replicas": replicas,
            "selector": {"matchLabels": {"statefulset": name}},       # :421-425
            "template": {"metadata": tmeta, "spec": pod_spec},
        },
    }


def generate_service(nb, env=None):
    """generateService, notebook_controller.go:486-513."""
    name, ns = _name_ns(nb)
    containers = _pod_spec(nb).get("containers")
    if not containers:
        raise ReferencePanic("index out of range: Containers[0]")     # :489
    port = DEFAULT_CONTAINER_PORT
    ports = containers[0].get("ports")
    if ports is not None:                                             # :490 (nil test, not len)
        if len(ports) == 0:
            raise ReferencePanic("index out of range: Ports[0]")      # :491
        port = int(ports[0].get("containerPort") or 0)
    return {
        "apiVersion": "v1", "kind": "Service",
        "metadata": _obj_meta(name, ns),
        "spec": {
            "ports": [{"name": "http-" + name, "protocol": "TCP",
                       "port": DEFAULT_SERVING_PORT, "targetPort": port}],   # :501-508
            "selector": {"statefulset": name},
            "type": "ClusterIP",
        },
    }


def virtual_service_name(name, ns):
    """virtualServiceName, notebook_controller.go:515-517."""
    return "notebook-%s-%s" % (ns, name)


def _headers_request_set(annotations):
    """notebook_controller.go:566-579: json.Unmarshal into map[string]string,
    empty map on any decode error.  Go semantics recalled: a JSON null value
    leaves the zero string; any other non-string value is an error."""
    raw = annotations.get(ANNOTATION_HEADERS_REQUEST_SET, "")
    if not raw:
        return {}
    try:
        v = json.loads(raw)
    except ValueError:
        return {}
    if v is None:
        return {}
    if not isinstance(v, dict):
        return {}
    out = {}
    for k, x in v.items():
        if x is None:
            out[k] = ""
        elif isinstance(x, str):
            out[k] = x
        else:
            return {}
    return out


def generate_virtualservice(nb, env):
    """generateVirtualService, notebook_controller.go:519-619."""
    name, ns = _name_ns(nb)
    prefix = "/notebook/%s/%s/" % (ns, name)                          # :523
    annotations = dict(get_map(meta(nb), "annotations"))
    rewrite = prefix                                                   # :531
    if annotations.get(ANNOTATION_REWRITE_URI):                        # :533
        rewrite = annotations[ANNOTATION_REWRITE_URI]
    cluster_domain = env["CLUSTER_DOMAIN"] if "CLUSTER_DOMAIN" in env else "cluster.local"  # :537
    service = "%s.%s.svc.%s" % (name, ns, cluster_domain)              # :540
    host = env.get("ISTIO_HOST") or "*"                                # :548-551
    gateway = env.get("ISTIO_GATEWAY") or "kubeflow/kubeflow-gateway"  # :557-560
    return {
        "apiVersion": "networking.istio.io/v1alpha3", "kind": "VirtualService",
        "metadata": _obj_meta(virtual_service_name(name, ns), ns),
        "spec": {
            "hosts": [host],
            "gateways": [gateway],
            "http": [{
                "headers": {"request": {"set": _headers_request_set(annotations)}},
                "match": [{"uri": {"prefix": prefix}}],
                "rewrite": {"uri": rewrite},
                "route": [{"destination": {"host": service,
                                           "port": {"number": DEFAULT_SERVING_PORT}}}],
            }],
        },
    }


# ---------------------------------------------------------------- reconcilehelper


def _copy_map_field(frm, to, field):
    """util.go:109-121 (and :168-180): iterate TO's keys only, then assign."""
    require = False
    fm, tm = meta(frm), to.setdefault("metadata", {})
    f = get_map(fm, field)
    for k, v in get_map(tm, field).items():
        if f.get(k, "") != v:
            require = True
    if fm.get(field) is None:
        tm.pop(field, None)
    else:
        tm[field] = deepcopy(fm[field])
    return require


def copy_statefulset_fields(frm, to):
    """reconcilehelper.CopyStatefulSetFields, util.go:107-134.  Mutates `to`."""
    require = _copy_map_field(frm, to, "labels")
    require |= _copy_map_field(frm, to, "annotations")
    fs, ts = frm.get("spec") or {}, to.setdefault("spec", {})
    if fs.get("replicas") is None or ts.get("replicas") is None:
        raise ReferencePanic("nil pointer dereference: Spec.Replicas")  # :123
    if fs["replicas"] != ts["replicas"]:
        ts["replicas"] = fs["replicas"]
        require = True
    ft = (fs.get("template") or {}).get("spec") or {}
    tt = ts.setdefault("template", {})
    if not json_equal(tt.get("spec") or {}, ft):                       # :128
        require = True
    tt["spec"] = deepcopy(ft)                                          # :131
    return require


def copy_service_fields(frm, to):
    """reconcilehelper.CopyServiceFields, util.go:166-195.  Mutates `to`."""
    require = _copy_map_field(frm, to, "labels")
    require |= _copy_map_field(frm, to, "annotations")
    fs, ts = frm.get("spec") or {}, to.setdefault("spec", {})
    for field in ("selector", "ports"):                                # :184-192
        if not json_equal(ts.get(field), fs.get(field)):
            require = True
        if fs.get(field) is None:
            ts.pop(field, None)
        else:
            ts[field] = deepcopy(fs[field])
    return require


def copy_virtualservice(frm, to):
    """reconcilehelper.CopyVirtualService, util.go:199-219.  Mutates `to`."""
    if not isinstance(frm.get("spec"), dict):                          # :200-206
        return False
    if not isinstance(to.get("spec"), dict):                           # :208-212
        to["spec"] = deepcopy(frm["spec"])
        return True
    require = not json_equal(frm["spec"], to["spec"])                  # :214
    if require:
        to["spec"] = deepcopy(frm["spec"])
    return require


def precopy_template_labels(desired, found):
    """notebook_controller.go:165-169: template labels follow only when the
    replica counts differ.  Mutates `found`."""
    dr = (desired.get("spec") or {}).get("replicas")
    fr = (found.get("spec") or {}).get("replicas")
    if dr is not None and fr is not None and dr != fr:
        dl = (((desired["spec"].get("template") or {}).get("metadata")) or {}).get("labels")
        ft = found["spec"].setdefault("template", {}).setdefault("metadata", {})
        if not json_equal(ft.get("labels"), dl):
            if dl is None:
                ft.pop("labels", None)
            else:
                ft["labels"] = deepcopy(dl)


def create_or_update(desired, found):
    """The create-or-update decision of Reconcile for one owned object
    (notebook_controller.go:147-179 StatefulSet, :189-210 Service, :636-657
    VirtualService).  found is None for NotFound.

    Returns the object the reference would send: `desired` itself (Create), the
    found object with the owned fields copied over (Update), or None."""
    if found is None:
        return deepcopy(desired)
    to = deepcopy(found)
    kind = desired.get("kind")
    if kind == "VirtualService":
        require = copy_virtualservice(desired, to)
    elif kind == "Service":
        require = copy_service_fields(desired, to)
    else:
        precopy_template_labels(desired, to)
        require = copy_statefulset_fields(desired, to)
    return to if require else None


# ---------------------------------------------------------------- status mirroring

_ZERO = (None, "", 0, False)


def _pod_status_is_zero(st):
    """reflect.DeepEqual(pod.Status, corev1.PodStatus{}) (:304): any non-nil
    slice, even an empty one, makes it non-zero."""
    if not st:
        return True
    return all((v in _ZERO and not isinstance(v, (list, dict))) for v in st.values())


def _state_is_zero(cs):
    return not cs or all(cs.get(k) is None for k in ("waiting", "running", "terminated"))


def pod_cond_to_notebook_cond(podc, now):
    """PodCondToNotebookCond, notebook_controller.go:351-390; `now` stands for
    metav1.Now() (:377,:386) as an RFC 3339 string."""
    c = {"type": podc.get("type") or "", "status": podc.get("status") or ""}
    c["lastProbeTime"] = podc.get("lastProbeTime") or now
    c["lastTransitionTime"] = podc.get("lastTransitionTime") or now
    if podc.get("reason"):
        c["reason"] = podc["reason"]
    if podc.get("message"):
        c["message"] = podc["message"]
    return c


def create_notebook_status(nb, sts, pod, now):
    """createNotebookStatus, notebook_controller.go:290-349.

    Golden: notebook_controller_test.go:103-275 (5 cases)."""
    status = {"conditions": [],
              "readyReplicas": int(((sts or {}).get("status") or {}).get("readyReplicas") or 0),
              "containerState": {}}
    pst = (pod or {}).get("status")
    if _pod_status_is_zero(pst):                                       # :304-307
        return status
    name = meta(nb).get("name", "")
    cur = (nb.get("status") or {}).get("containerState") or {}
    for cs in pst.get("containerStatuses") or []:                      # :314-330
        if cs.get("name", "") != name:
            continue
        st = cs.get("state") or {}
        # Go `==` on ContainerState compares its three pointers: equal only
        # when both sides are all-nil (:319).
        if _state_is_zero(st) and _state_is_zero(cur):
            continue
        status["containerState"] = deepcopy(st)
        break
    status["conditions"] = [pod_cond_to_notebook_cond(c, now)
                            for c in pst.get("conditions") or []]      # :339-346
    return status


# ---------------------------------------------------------------- event plumbing


def nb_name_from_involved_object(get_pod, ref):
    """nbNameFromInvolvedObject, notebook_controller.go:666-690.
    get_pod(namespace, name) -> pod tree or None.  Returns (name, err)."""
    if ref.get("kind") == "StatefulSet":
        return ref.get("name", ""), None
    if ref.get("kind") == "Pod":
        pod = get_pod(ref.get("namespace", ""), ref.get("name", ""))
        if pod is None:
            return "", "not found"
        labels = get_map(meta(pod), "labels")
        if "notebook-name" in labels:
            return labels["notebook-name"], None
    return "", "object isn't related to a Notebook"


def is_sts_or_pod_event(event):
    """isStsOrPodEvent, notebook_controller.go:662-664."""
    return (event.get("involvedObject") or {}).get("kind") in ("Pod", "StatefulSet")


# ---------------------------------------------------------------- version conversion
API_VERSIONS = ("kubeflow.org/v1beta1", "kubeflow.org/v1", "kubeflow.org/v1alpha1")


def convert_notebook(src, to_version, keep_metadata=False):
    """ConvertTo / ConvertFrom, components/notebook-controller/api/v1/notebook_conversion.go:25-69
    (api/v1alpha1 has the same two functions): the hub is v1beta1, a spoke converts to it with
    ConvertTo and from it with ConvertFrom, and both copy exactly four things —
    spec.template.spec, status.readyReplicas, status.containerState, and of every condition its
    type, lastProbeTime, reason and message.  A condition's `status` and `lastTransitionTime`
    are NOT copied (:30-38, :56-64): the destination has status "" and a zero time.  ObjectMeta
    is not touched by these functions; keep_metadata carries the source's through, as a caller
    that wants a usable object has to."""
    spec = ((src.get("spec") or {}).get("template") or {}).get("spec")
    st = src.get("status") or {}
    conds = []
    for c in st.get("conditions") or []:
        n = {"type": c.get("type", ""), "status": ""}
        if c.get("lastProbeTime"):
            n["lastProbeTime"] = c["lastProbeTime"]
        if c.get("reason"):
            n["reason"] = c["reason"]
        if c.get("message"):
            n["message"] = c["message"]
        conds.append(n)
    out = {"apiVersion": to_version, "kind": "Notebook", "metadata": {},
           "spec": {"template": {"spec": deepcopy(spec) if isinstance(spec, dict) else {}}},
           "status": {"conditions": conds, "readyReplicas": st.get("readyReplicas") or 0,
                      "containerState": deepcopy(st.get("containerState")) if isinstance(st.get("containerState"), dict) else {}}}
    if keep_metadata and isinstance(src.get("metadata"), dict):
        out["metadata"] = deepcopy(src["metadata"])
    return out
