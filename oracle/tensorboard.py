"""Oracle (test infrastructure only): the desired objects of one Tensorboard reconcile.

Follows components/tensorboard-controller/controllers/tensorboard_controller.go:
generateDeployment :167-299, generateService :301-319, generateVirtualService :321-378, the path
helpers :380-412 (isCloudPath, isGoogleCloudPath, isPVCPath, extractPVCName, extractPVCSubPath),
generateNodeAffinity's result :445-465, rwoPVCScheduling :471-482, CopyDeploymentSetFields
:512-533, the status update of Reconcile :119-152, and ctrl.SetControllerReference at :94,:103,
:112.  The PVC Get and the Pod List of the RWO scheduling path are I/O: the node name they lead to
is a parameter.  The reference has no test that asserts these objects (controllers/suite_test.go is
the envtest bootstrap only): parity unpinned, the restatement is anchored on the Go source.
"""
from .gojson import deepcopy, meta

ENV_REQUIRED = ("TENSORBOARD_IMAGE", "ISTIO_GATEWAY", "ISTIO_HOST")


class ReferenceError_(Exception):
    """The reference returns an error here (a required environment variable is not set, or
    RWO_PVC_SCHEDULING has an invalid value)."""


def is_google_cloud_path(p):
    return p.startswith("gs://")                                                        # :384-386


def is_cloud_path(p):
    return is_google_cloud_path(p) or p.startswith("s3://") or p.startswith("/cns/")    # :380-382


def is_pvc_path(p):
    return p.startswith("pvc://")                                                       # :388-390


def extract_pvc_name(p):
    t = p[len("pvc://"):] if p.startswith("pvc://") else p                              # :392-400
    e = t.find("/")
    return t if e == -1 else t[:e]


def extract_pvc_subpath(p):
    t = p[len("pvc://"):] if p.startswith("pvc://") else p                              # :402-410
    s = t.find("/")
    return "" if s == -1 or len(t) == s + 1 else t[s + 1:]


def rwo_pvc_scheduling(env):
    """rwoPVCScheduling :471-482."""
    if "RWO_PVC_SCHEDULING" not in env or env["RWO_PVC_SCHEDULING"] in ("false", "False", "FALSE"):
        return False
    if env["RWO_PVC_SCHEDULING"] in ("true", "True", "TRUE"):
        return True
    raise ReferenceError_("Invalid value for 'RWO_PVC_SCEDULING' env var.")


def _getenv(env, k):
    if k not in env:                                                                    # :535-541 (empty counts as set)
        raise ReferenceError_("environment variable %s is not set" % k)
    return env[k]


def owner_ref(tb):
    m = meta(tb)
    return {"apiVersion": "tensorboard.kubeflow.org/v1alpha1", "kind": "Tensorboard", "name": m.get("name", ""),
            "uid": m.get("uid", ""), "controller": True, "blockOwnerDeletion": True}


def _obj_meta(tb, owner):
    m = {"name": meta(tb).get("name", "")}
    if meta(tb).get("namespace", ""):
        m["namespace"] = meta(tb)["namespace"]
    if owner:
        m["ownerReferences"] = [owner_ref(tb)]
    return m


def node_affinity(nodename):
    """What generateNodeAffinity leaves in *affinity :445-465; {} when there is no node."""
    if not nodename:
        return {}
    return {"nodeAffinity": {"preferredDuringSchedulingIgnoredDuringExecution": [
        {"weight": 100, "preference": {"matchExpressions": [
            {"key": "kubernetes.io/hostname", "operator": "In", "values": [nodename]}]}}]}}


def generate_deployment(tb, env, nodename="", owner=True):
    """generateDeployment :167-299.  nodename: the node of a running Pod that mounts the PVC
    ReadWriteOnce, as the RWO scheduling path finds it ("" = none, or scheduling off)."""
    name = meta(tb).get("name", "")
    logspath = (tb.get("spec") or {}).get("logspath", "") or ""
    image = _getenv(env, "TENSORBOARD_IMAGE")                                           # :172-175
    mountpath, subpath = logspath, ""
    mounts, volumes = [], []
    affinity = {}
    if not is_cloud_path(logspath):                                                     # :178
        if is_pvc_path(logspath):                                                       # :182-185
            pvcname = extract_pvc_name(logspath)
            mountpath = "/tensorboard_logs/"
            subpath = extract_pvc_subpath(logspath)
        else:
            pvcname, subpath = "tb-volume", ""                                          # :188-189
        m = {"name": "tbpd", "readOnly": True, "mountPath": mountpath}
        if subpath:
            m["subPath"] = subpath                                                      # omitempty
        mounts.append(m)
        volumes.append({"name": "tbpd", "persistentVolumeClaim": {"claimName": pvcname}})
        if rwo_pvc_scheduling(env):                                                     # :207-232
            affinity = node_affinity(nodename)
    elif is_google_cloud_path(logspath):                                                # :233-249
        mounts.append({"name": "gcp-creds", "readOnly": True, "mountPath": "/secret/gcp"})
        volumes.append({"name": "gcp-creds", "secret": {"secretName": "user-gcp-sa"}})
    labels = dict(meta(tb).get("labels") or {})                                         # :252-256
    labels["app"] = name
    container = {"name": "tensorboard", "image": image, "imagePullPolicy": "IfNotPresent",
                 "command": ["/usr/local/bin/tensorboard"], "workingDir": "/",
                 "args": ["--logdir=" + mountpath, "--bind_all"], "ports": [{"containerPort": 6006}]}
    if not image:
        del container["image"]                                                          # omitempty
    if mounts:
        container["volumeMounts"] = mounts
    pod = {"affinity": affinity, "restartPolicy": "Always", "containers": [container]}
    if volumes:
        pod["volumes"] = volumes
    return {"apiVersion": "apps/v1", "kind": "Deployment", "metadata": _obj_meta(tb, owner),
            "spec": {"replicas": 1, "selector": {"matchLabels": {"app": name}},
                     "template": {"metadata": {"labels": labels}, "spec": pod}}}


def generate_service(tb, owner=True):
    """generateService :301-319."""
    name = meta(tb).get("name", "")
    return {"apiVersion": "v1", "kind": "Service", "metadata": _obj_meta(tb, owner),
            "spec": {"type": "ClusterIP", "selector": {"app": name},
                     "ports": [{"name": "http-" + name, "port": 80, "targetPort": 6006}]}}


def generate_virtualservice(tb, env, owner=True):
    """generateVirtualService :321-378."""
    name, ns = meta(tb).get("name", ""), meta(tb).get("namespace", "")
    gateway = _getenv(env, "ISTIO_GATEWAY")
    host = _getenv(env, "ISTIO_HOST")
    m = {"name": name}
    if ns:
        m["namespace"] = ns                       # SetNamespace("") removes the field
    if owner:
        m["ownerReferences"] = [owner_ref(tb)]
    return {"apiVersion": "networking.istio.io/v1alpha3", "kind": "VirtualService", "metadata": m,
            "spec": {"hosts": [host], "gateways": [gateway],
                     "http": [{"match": [{"uri": {"prefix": "/tensorboard/%s/%s/" % (ns, name)}}],
                               "rewrite": {"uri": "/"},
                               "route": [{"destination": {"host": "%s.%s.svc.cluster.local" % (name, ns),
                                                          "port": {"number": 80}}}],
                               "timeout": "300s"}]}}


def tensorboard_objects(tb, env, nodename="", owner=True):
    """(Deployment, Service, VirtualService) of one reconcile, in the order Reconcile builds them."""
    for k in ENV_REQUIRED:
        _getenv(env, k)
    return (generate_deployment(tb, env, nodename, owner), generate_service(tb, owner),
            generate_virtualservice(tb, env, owner))


def copy_deployment_fields(frm, to):
    """CopyDeploymentSetFields :512-533: (updated `to`, requireUpdate)."""
    to = deepcopy(to)
    update = False
    to_labels = meta(to).get("labels") or {}
    from_labels = meta(frm).get("labels") or {}
    for k, v in to_labels.items():                                                      # :514-518
        if from_labels.get(k, "") != v:
            update = True
    to.setdefault("metadata", {})
    if meta(frm).get("labels") is None:                                                 # :519 to.Labels = from.Labels
        to["metadata"].pop("labels", None)
    else:
        to["metadata"]["labels"] = deepcopy(meta(frm)["labels"])
    fr, tr = (frm.get("spec") or {}).get("replicas"), (to.get("spec") or {}).get("replicas")
    if fr is None or tr is None:
        raise ReferenceError_("nil pointer dereference: Spec.Replicas")                 # :521
    if fr != tr:
        to["spec"]["replicas"] = fr
        update = True

    def aff(o):
        return ((((o.get("spec") or {}).get("template") or {}).get("spec")) or {}).get("affinity")
    if aff(to) != aff(frm):                                                             # :526-529 (text DeepEqual)
        update = True
    tspec = to.setdefault("spec", {}).setdefault("template", {}).setdefault("spec", {})
    if aff(frm) is None:
        tspec.pop("affinity", None)
    else:
        tspec["affinity"] = deepcopy(aff(frm))
    return to, update


def update_status(tb, found_deployment):
    """Reconcile :119-152: the Tensorboard's status after looking at the found Deployment
    (None = NotFound: nothing is written).  Returns the new status or None."""
    if found_deployment is None:
        return None
    status = deepcopy(tb.get("status") or {})
    conds = list(status.get("conditions") or [])
    dconds = (found_deployment.get("status") or {}).get("conditions") or []
    if dconds:                                                                          # :131-141
        # metav1.Time has no omitempty that works on a struct: the zero time is written as null
        c = {"deploymentState": dconds[0].get("type", ""), "lastProbeTime": dconds[0].get("lastUpdateTime")}
        if not conds or conds[-1].get("deploymentState", "") != c["deploymentState"]:
            conds.append(c)
    status["conditions"] = conds if conds else None      # `json:"conditions"`: a nil slice is null
    status["readyReplicas"] = (found_deployment.get("status") or {}).get("readyReplicas", 0)   # :143-146
    return status
