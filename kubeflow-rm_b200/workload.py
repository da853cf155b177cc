"""Synthetic inputs for the parity tests and bench.py.

All generators are deterministic in (seed, index).  Shapes follow the
reference's own fixtures:
  C1  the envtest Notebook of notebook-controller/controllers/
      notebook_controller_bdd_test.go:46-57 (one busybox container)
  C2  the OAuth Notebook of odh-notebook-controller/controllers/
      notebook_controller_test.go:573-600
  loadtest  notebook-controller/loadtest/jupyter_test.yaml
Names are nb-%06d, namespaces ns-%03d with 100 Notebooks per namespace.
"""
import json
import random

SEED = 20261106


def _meta(i, extra=None):
    m = {"name": "nb-%06d" % i, "namespace": "ns-%03d" % (i // 100),
         "uid": "%08x-0000-4000-8000-%012x" % (i & 0xffffffff, i),
         "resourceVersion": str(1000 + i), "generation": 1,
         "creationTimestamp": "2026-11-06T00:00:00Z"}
    if extra:
        m.update(extra)
    return m


def notebook_c1(i=0):
    """C1: the envtest object (name/namespace varied by index)."""
    if i == 0:
        md = {"name": "test-notebook", "namespace": "default"}
    else:
        md = _meta(i)
    return {"apiVersion": "kubeflow.org/v1beta1", "kind": "Notebook", "metadata": md,
            "spec": {"template": {"spec": {"containers": [{"name": "busybox", "image": "busybox"}]}}}}


def notebook_c2(i):
    """C2: the ODH OAuth fixture, one per index."""
    md = _meta(i)
    name, ns = md["name"], md["namespace"]
    md["labels"] = {"app.kubernetes.io/instance": name}
    md["annotations"] = {
        "notebooks.opendatahub.io/inject-oauth": "true",
        "notebooks.opendatahub.io/foo": "bar",
        "notebooks.opendatahub.io/oauth-logout-url":
            "https://example.notebook-url/notebook/" + ns + "/" + name,
    }
    return {"apiVersion": "kubeflow.org/v1", "kind": "Notebook", "metadata": md,
            "spec": {"template": {"spec": {
                "containers": [{"name": name, "image": "registry.redhat.io/ubi8/ubi:latest"}],
                "volumes": [{"name": "notebook-data",
                             "persistentVolumeClaim": {"claimName": name + "-data"}}]}}}}


def notebook_loadtest(i):
    """notebook-controller/loadtest/jupyter_test.yaml with start_notebooks.py's renaming."""
    md = _meta(i)
    return {"apiVersion": "kubeflow.org/v1", "kind": "Notebook", "metadata": md,
            "spec": {"template": {"spec": {
                "serviceAccountName": "default-editor",
                "volumes": [{"name": "test-vol", "persistentVolumeClaim": {"claimName": "test-pvc-%d" % i}},
                            {"emptyDir": {"medium": "Memory"}, "name": "dshm"}],
                "containers": [{
                    "image": "gcr.io/kubeflow-images-public/tensorflow-1.14.0-notebook-cpu:v1.0.0",
                    "name": "notebook",
                    "resources": {"requests": {"cpu": "500m", "memory": "1Gi"}},
                    "volumeMounts": [{"mountPath": "/home/jovyan", "name": "test-vol"},
                                     {"mountPath": "/dev/shm", "name": "dshm"}]}]}}}}


_LABEL_KEYS = ["app", "app.kubernetes.io/instance", "opendatahub.io/dashboard", "team", "tier",
               "statefulset", "notebook-name", "opendatahub.io/workbenches", "access-ml-pipeline",
               "opendatahub.io/odh-managed"]
_ANN_KEYS = ["foo", "kubectl.kubernetes.io/last-applied-configuration",
             "notebooks.opendatahub.io/inject-oauth", "opendatahub.io/username",
             "openshift.io/display-name", "openshift.io/description",
             "notebooks.kubeflow.org/last-activity", "kubeflow-resource-stopped",
             "opendatahub.io/image-display-name", "notebooks.opendatahub.io/last-size-selection"]
_WEIRD = ["", "x", "a\"b", "back\\slash", "uni\u00e9\u4e2d", "tab\there", "new\nline", "<&>",
          "{\"k\":\"v\"}", "/some/path/", "sp ace", "\u0001ctl"]


def notebook_fuzz(rng, i):
    """A randomly shaped Notebook that walks every branch of the generators."""
    nb = notebook_c2(i) if rng.random() < 0.5 else notebook_loadtest(i)
    md = nb["metadata"]
    if rng.random() < 0.05:
        md.pop("namespace")
    if rng.random() < 0.1:
        md.pop("uid")
    r = rng.random()
    if r < 0.15:
        md.pop("labels", None)
    elif r < 0.25:
        md["labels"] = {}
    else:
        md["labels"] = {k: rng.choice(_WEIRD[1:4] + ["true", "v1"])
                        for k in rng.sample(_LABEL_KEYS, rng.randint(1, 6))}
    r = rng.random()
    if r < 0.15:
        md.pop("annotations", None)
    elif r < 0.25:
        md["annotations"] = {}
    else:
        md["annotations"] = {k: rng.choice(_WEIRD + ["true", "2026-11-06T10:00:00Z"])
                             for k in rng.sample(_ANN_KEYS, rng.randint(1, 6))}
        if rng.random() < 0.4:
            md["annotations"]["notebooks.kubeflow.org/http-rewrite-uri"] = rng.choice(["", "/", "/lab/x"])
        if rng.random() < 0.5:
            md["annotations"]["notebooks.kubeflow.org/http-headers-request-set"] = rng.choice([
                "", "{}", "{\"X-RStudio-Root-Path\":\"/notebook/ns/nb/\"}",
                "{\"A\":\"b\\\"q\",\"B\":null, \"C\" : \"\\u00e9\"}", "{\"A\":1}", "[1,2]", "null",
                "not json", "{\"A\":\"b\"} trailing", " { \"A\" : \"b\" } ", "{\"A\":{\"n\":1}}",
                "{\"A\":\"b\",}", "\"str\""])
    if md.get("annotations") is not None:
        a = md["annotations"]
        if rng.random() < 0.5:
            a["notebooks.opendatahub.io/inject-oauth"] = rng.choice(
                ["true", "false", "1", "T", "TRUE", "True", "t", "yes", "", "0", "tRuE"])
        if rng.random() < 0.2:
            a["opendatahub.io/service-mesh"] = rng.choice(["true", "false", "1", "nope"])
        if rng.random() < 0.3:
            a["notebooks.opendatahub.io/oauth-logout-url"] = rng.choice(
                ["", "https://x.example/logout?a=b&c=\"d\"", "https://example/notebook/ns/nb"])
    spec = nb["spec"]["template"]["spec"]
    c0 = spec["containers"][0]
    r = rng.random()
    if r < 0.2:
        c0["workingDir"] = "/opt/app-root/src"
    elif r < 0.3:
        c0["workingDir"] = ""
    r = rng.random()
    if r < 0.25:
        c0["ports"] = [{"containerPort": rng.choice([8888, 8080, 8787]), "name": "nb", "protocol": "TCP"}]
    elif r < 0.3:
        c0["ports"] = None
    elif r < 0.33:
        c0["ports"] = []
    elif r < 0.36:
        c0["ports"] = [{"name": "noport"}]
    r = rng.random()
    if r < 0.3:
        c0["env"] = [{"name": "JUPYTER_NOTEBOOK_PORT", "value": "8888"},
                     {"name": "NOTEBOOK_ARGS", "value": "--NotebookApp.token='' --x=\"y\""}]
    elif r < 0.4:
        c0["env"] = [{"name": "A", "value": "1"}, {"name": "NB_PREFIX", "value": "/custom"}]
    elif r < 0.45:
        c0["env"] = []
    elif r < 0.5:
        c0["env"] = None
    r = rng.random()
    if r < 0.15:
        spec["securityContext"] = {"fsGroup": 1001, "runAsUser": 1001}
    elif r < 0.2:
        spec["securityContext"] = {}
    elif r < 0.25:
        spec["securityContext"] = None
    if rng.random() < 0.3:
        spec["containers"].append({"name": "sidecar", "image": "side:1", "args": ["--x", "[]{}\","]})
    if rng.random() < 0.1:
        spec["containers"] = []
    if rng.random() < 0.03:
        nb["spec"] = {}
    if rng.random() < 0.2:
        nb["status"] = {"conditions": [], "readyReplicas": 1, "containerState": {}}
    return nb


def dumps(tree, rng=None):
    """JSON bytes; with an rng, sometimes pretty-printed or ASCII-escaped so the
    walker sees whitespace and \\u escapes."""
    if rng is not None:
        r = rng.random()
        if r < 0.15:
            return json.dumps(tree, indent=2, ensure_ascii=False).encode()
        if r < 0.3:
            return json.dumps(tree, separators=(", ", ": "), ensure_ascii=True).encode()
    return json.dumps(tree, separators=(",", ":"), ensure_ascii=False).encode()


def corpus(kind, n, seed=SEED, start=0):
    """n Notebook trees of one kind: 'c1', 'c2', 'loadtest' or 'fuzz'."""
    if kind == "c1":
        return [notebook_c1(start + i) for i in range(n)]
    if kind == "c2":
        return [notebook_c2(start + i) for i in range(n)]
    if kind == "loadtest":
        return [notebook_loadtest(start + i) for i in range(n)]
    if kind == "fuzz":
        rng = random.Random(seed + start)
        return [notebook_fuzz(rng, start + i) for i in range(n)]
    raise ValueError(kind)


def found_variant(rng, desired):
    """A "found" object for `desired` as an API server might return it: None
    (NotFound), the same object with server-side metadata and status, or a copy
    that differs in an owned or an unowned field.  Used by the Copy*Fields
    parity tests (reference: common/reconcilehelper/util.go:107-219)."""
    import copy
    r = rng.random()
    if r < 0.12:
        return None
    f = copy.deepcopy(desired)
    md = f.setdefault("metadata", {})
    md.update({"uid": "f-%d" % rng.randrange(10**6), "resourceVersion": str(rng.randrange(10**6)),
               "creationTimestamp": "2026-11-06T00:00:00Z"})
    kind = desired.get("kind")
    if kind == "StatefulSet":
        f["status"] = {"replicas": 1, "readyReplicas": rng.randint(0, 1)}
        f["spec"]["serviceName"] = ""
        f["spec"]["podManagementPolicy"] = "OrderedReady"
    elif kind == "Service":
        f["spec"]["clusterIP"] = "10.0.%d.%d" % (rng.randrange(256), rng.randrange(256))
        f["spec"]["sessionAffinity"] = "None"
        f["status"] = {"loadBalancer": {}}
    r = rng.random()
    if r < 0.35:
        pass                                             # unchanged: no update expected
    elif r < 0.45:
        md["labels"] = {"added-by": rng.choice(["someone", ""])}
    elif r < 0.5:
        md["annotations"] = {"note": "x", "empty": ""}
    elif r < 0.6 and kind == "StatefulSet":
        f["spec"]["replicas"] = 1 - f["spec"]["replicas"]
        if rng.random() < 0.5:
            f["spec"]["template"]["metadata"]["labels"]["extra"] = "1"
    elif r < 0.75 and kind == "StatefulSet":
        ps = f["spec"]["template"]["spec"]
        ps["dnsPolicy"] = "ClusterFirst"                 # server-side defaulting
        ps["containers"][0]["terminationMessagePath"] = "/dev/termination-log"
    elif r < 0.8 and kind == "StatefulSet":
        f["spec"]["template"]["spec"]["containers"][0]["image"] = "other:latest"
    elif r < 0.85 and kind == "StatefulSet":
        f["spec"]["template"]["metadata"]["labels"]["drift"] = "1"   # not owned unless replicas differ
    elif r < 0.7 and kind == "Service":
        f["spec"]["ports"][0]["targetPort"] = 9999
    elif r < 0.8 and kind == "Service":
        f["spec"]["selector"] = {"statefulset": "someone-else"}
    elif r < 0.85 and kind == "Service":
        f["spec"]["ports"][0]["nodePort"] = None          # null member = nil: still equal
    elif r < 0.7 and kind == "VirtualService":
        f["spec"]["hosts"] = ["other.example.com"]
    elif r < 0.8 and kind == "VirtualService":
        f["spec"]["http"][0]["timeout"] = "300s"
    elif r < 0.85 and kind == "VirtualService":
        del f["spec"]
    if rng.random() < 0.3:                                # key order must not matter
        def shuffle(x):
            if isinstance(x, dict):
                ks = list(x)
                rng.shuffle(ks)
                return {k: shuffle(x[k]) for k in ks}
            if isinstance(x, list):
                return [shuffle(v) for v in x]
            return x
        f = shuffle(f)
    return f


# ---------------------------------------------------------------- API-server-shaped objects
# What a Get returns after the object went through an API server: members in the order of
# the Go structs' fields (k8s.io/api, recalled from v0.29 — nothing asserts it here, and the
# library does not depend on it: DeepEqual is order-free), the fields defaulting adds to a
# StatefulSet / Service, server-side metadata with managedFields, and a status.
_GO_ORDER = {
    "": ["kind", "apiVersion", "metadata", "spec", "status"],
    "metadata": ["name", "generateName", "namespace", "uid", "resourceVersion", "generation", "creationTimestamp",
                 "deletionTimestamp", "labels", "annotations", "ownerReferences", "finalizers", "managedFields"],
    "sts.spec": ["replicas", "selector", "template", "volumeClaimTemplates", "serviceName", "podManagementPolicy",
                 "updateStrategy", "revisionHistoryLimit", "minReadySeconds", "persistentVolumeClaimRetentionPolicy"],
    "podspec": ["volumes", "initContainers", "containers", "ephemeralContainers", "restartPolicy",
                "terminationGracePeriodSeconds", "activeDeadlineSeconds", "dnsPolicy", "nodeSelector",
                "serviceAccountName", "serviceAccount", "automountServiceAccountToken", "nodeName", "hostNetwork",
                "hostPID", "hostIPC", "shareProcessNamespace", "securityContext", "imagePullSecrets", "hostname",
                "subdomain", "affinity", "schedulerName", "tolerations", "hostAliases", "priorityClassName", "priority"],
    "container": ["name", "image", "command", "args", "workingDir", "ports", "envFrom", "env", "resources",
                  "resizePolicy", "restartPolicy", "volumeMounts", "volumeDevices", "livenessProbe", "readinessProbe",
                  "startupProbe", "lifecycle", "terminationMessagePath", "terminationMessagePolicy", "imagePullPolicy",
                  "securityContext", "stdin", "stdinOnce", "tty"],
    "svc.spec": ["ports", "selector", "clusterIP", "clusterIPs", "type", "externalIPs", "sessionAffinity",
                 "loadBalancerIP", "ipFamilies", "ipFamilyPolicy", "internalTrafficPolicy"],
}


def _ordered(d, which):
    first = [k for k in _GO_ORDER[which] if k in d]
    return {k: d[k] for k in first + [k for k in d if k not in first]}


def server_shaped(desired, i=0):
    """`desired` (StatefulSet, Service or an unstructured object such as a VirtualService)
    as the API server would hand it back on the next Get."""
    import copy
    f = copy.deepcopy(desired)
    kind = f.get("kind")
    md = f.setdefault("metadata", {})
    md.update({"uid": "%08x-aaaa-4bbb-8ccc-%012x" % (i & 0xffffffff, i), "resourceVersion": str(70000 + i),
               "generation": 1, "creationTimestamp": "2026-11-06T00:00:00Z",
               "managedFields": [{"manager": "manager", "operation": "Update", "apiVersion": f.get("apiVersion"),
                                  "time": "2026-11-06T00:00:00Z", "fieldsType": "FieldsV1",
                                  "fieldsV1": {"f:metadata": {"f:ownerReferences": {".": {}}}, "f:spec": {}}}]})
    f["metadata"] = _ordered(md, "metadata")
    if kind == "StatefulSet":
        sp = f["spec"]
        sp.update({"serviceName": "", "podManagementPolicy": "OrderedReady",
                   "updateStrategy": {"type": "RollingUpdate", "rollingUpdate": {"partition": 0}},
                   "revisionHistoryLimit": 10,
                   "persistentVolumeClaimRetentionPolicy": {"whenDeleted": "Retain", "whenScaled": "Retain"}})
        t = sp["template"]
        t["metadata"] = _ordered(dict(t.get("metadata") or {}, creationTimestamp=None), "metadata")
        ps = t["spec"]
        ps.update({"restartPolicy": "Always", "terminationGracePeriodSeconds": 30, "dnsPolicy": "ClusterFirst",
                   "schedulerName": "default-scheduler"})
        ps.setdefault("securityContext", {})
        if ps.get("serviceAccountName"):
            ps["serviceAccount"] = ps["serviceAccountName"]
        for c in ps.get("containers") or []:
            c.setdefault("resources", {})
            c.update({"terminationMessagePath": "/dev/termination-log", "terminationMessagePolicy": "File",
                      "imagePullPolicy": "Always" if str(c.get("image", "")).endswith(":latest") or ":" not in str(c.get("image", "")) else "IfNotPresent"})
            for p in c.get("ports") or []:
                if isinstance(p, dict):
                    p.setdefault("protocol", "TCP")
        ps["containers"] = [_ordered(c, "container") for c in ps.get("containers") or []]
        for v in ps.get("volumes") or []:
            if isinstance(v.get("secret"), dict):
                v["secret"].setdefault("defaultMode", 420)
        t["spec"] = _ordered(ps, "podspec")
        f["spec"] = _ordered(sp, "sts.spec")
        f["status"] = {"observedGeneration": 1, "replicas": 1, "readyReplicas": 1, "currentReplicas": 1,
                       "updatedReplicas": 1, "currentRevision": "%s-5d8c9f7b6" % md.get("name", ""),
                       "updateRevision": "%s-5d8c9f7b6" % md.get("name", ""), "collisionCount": 0, "availableReplicas": 1}
    elif kind == "Service":
        sp = f["spec"]
        ip = "10.%d.%d.%d" % ((i >> 16) & 255, (i >> 8) & 255, i & 255)
        sp.update({"clusterIP": ip, "clusterIPs": [ip], "sessionAffinity": "None", "ipFamilies": ["IPv4"],
                   "ipFamilyPolicy": "SingleStack", "internalTrafficPolicy": "Cluster"})
        f["spec"] = _ordered(sp, "svc.spec")
        f["status"] = {"loadBalancer": {}}
    else:  # a custom resource: stored as it was sent, Go maps print their keys sorted
        def srt(x):
            if isinstance(x, dict):
                return {k: srt(x[k]) for k in sorted(x)}
            if isinstance(x, list):
                return [srt(v) for v in x]
            return x
        if "spec" in f:
            f["spec"] = srt(f["spec"])
    return _ordered(f, "")


def _stamp(t):
    import datetime
    return datetime.datetime.fromtimestamp(t, datetime.timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ")



def odh_found_variant(rng, desired):
    """A "found" ODH object for `desired` as the API server might return it: None
    (NotFound), the same object with server-side fields (a Route gets its host from the
    ingress controller), or one that drifted in labels, spec or subjects.  Used by the
    ODH write-decision parity tests (odh-notebook-controller/controllers/
    notebook_route.go:76-139, notebook_network.go:66-122, notebook_rbac.go:89-140)."""
    import copy
    if rng.random() < 0.12:
        return None
    f = copy.deepcopy(desired)
    md = f.setdefault("metadata", {})
    md.update({"uid": "o-%d" % rng.randrange(10**6), "resourceVersion": str(rng.randrange(10**6)),
               "creationTimestamp": "2026-11-06T00:00:00Z"})
    kind = desired.get("kind")
    if kind == "Route":
        f.setdefault("spec", {})["host"] = "nb-%d.apps.example.com" % rng.randrange(10**4)
        f["status"] = {"ingress": [{"host": f["spec"]["host"], "routerName": "default"}]}
    r = rng.random()
    if r < 0.45:
        pass
    elif r < 0.55:
        md["labels"] = dict(md.get("labels") or {}, drift="1")
    elif r < 0.6:
        md.pop("labels", None)
    elif r < 0.65:
        md["annotations"] = {"someone": "wrote this"}          # not compared
    elif r < 0.8 and isinstance(f.get("spec"), dict) and f["spec"]:
        k = rng.choice(sorted(f["spec"]))
        if rng.random() < 0.5:
            del f["spec"][k]
        else:
            f["spec"][k] = {"changed": [1, "two", None]}
    elif r < 0.85 and isinstance(f.get("spec"), dict):
        f["spec"]["addedByServer"] = {"x": 1}
    elif r < 0.9 and "subjects" in f:
        f["subjects"] = f["subjects"] + [{"kind": "User", "name": "eve"}]
    elif r < 0.93 and "spec" in f:
        del f["spec"]
    return f


def pod_variant(rng, nb, now=1793966400):
    """A Pod "<name>-0" for the status-mirroring tests, or None (NotFound).
    Shapes after notebook-controller/controllers/notebook_controller_test.go:103-275."""
    r = rng.random()
    if r < 0.12:
        return None
    name = (nb.get("metadata") or {}).get("name", "")
    pod = {"apiVersion": "v1", "kind": "Pod",
           "metadata": {"name": name + "-0", "namespace": (nb.get("metadata") or {}).get("namespace", "")}}
    if r < 0.2:
        return pod                                    # no status at all
    if r < 0.26:
        pod["status"] = {}
        return pod
    if r < 0.3:
        pod["status"] = {"phase": "", "hostIP": ""}   # all-zero scalars: still the zero PodStatus
        return pod
    st = {"phase": rng.choice(["Pending", "Running"])}
    if rng.random() < 0.85:
        conds = []
        for ty in rng.sample(["Initialized", "Ready", "ContainersReady", "PodScheduled"], rng.randint(0, 4)):
            c = {"type": ty, "status": rng.choice(["True", "False", "false"])}
            if rng.random() < 0.7:
                c["lastProbeTime"] = rng.choice([None, _stamp(now - rng.randrange(5000))])
            if rng.random() < 0.8:
                c["lastTransitionTime"] = _stamp(now - rng.randrange(5000))
            if rng.random() < 0.4:
                c["reason"] = rng.choice(["Unschedulable", "ContainersNotReady", ""])
            if rng.random() < 0.4:
                c["message"] = rng.choice(["0/1 nodes are available: 1 Insufficient cpu.",
                                           "containers with unready status: [\"nb\"]", ""])
            conds.append(c)
        st["conditions"] = conds
    if rng.random() < 0.85:
        css = []
        for cname in rng.sample([name, "oauth-proxy", "istio-proxy"], rng.randint(1, 3)):
            state = rng.choice([{"running": {"startedAt": _stamp(now - 100)}},
                                {"waiting": {"reason": "PodInitializing"}},
                                {"terminated": {"exitCode": 1, "reason": "Error"}},
                                {"running": {"startedAt": None}}, {}, None])
            cs = {"name": cname, "ready": rng.random() < 0.5, "restartCount": rng.randrange(3),
                  "image": "img", "imageID": ""}
            if state is not None or rng.random() < 0.5:
                cs["state"] = state
            css.append(cs)
        st["containerStatuses"] = css
    pod["status"] = st
    return pod


def sts_variant(rng, nb):
    if rng.random() < 0.15:
        return None
    s = {"apiVersion": "apps/v1", "kind": "StatefulSet", "metadata": dict(nb.get("metadata") or {})}
    r = rng.random()
    if r < 0.6:
        s["status"] = {"replicas": 1, "readyReplicas": rng.randint(0, 1)}
    elif r < 0.8:
        s["status"] = {"replicas": 1}
    return s


def status_notebook(rng, nb):
    """The Notebook as stored, sometimes with a previous status."""
    import copy
    nb = copy.deepcopy(nb)
    r = rng.random()
    if r < 0.3:
        nb["status"] = {"conditions": [], "readyReplicas": 0, "containerState": {}}
    elif r < 0.5:
        nb["status"] = {"conditions": [], "readyReplicas": 1,
                        "containerState": {"running": {"startedAt": "2026-11-06T00:00:00Z"}}}
    return nb


def culling_case(rng, i, now):
    """(notebook, pod_exists, kernels, terminals) for one culling pass; kernels /
    terminals are the decoded /api/kernels, /api/terminals bodies or None when
    the GET failed (culling_controller.go:209-279)."""
    nb = notebook_loadtest(i)
    a = {}
    r = rng.random()
    if r < 0.15:
        a["kubeflow-resource-stopped"] = rng.choice([_stamp(now - 50), "odh-notebook-controller-lock"])
    if rng.random() < 0.75:
        a["notebooks.kubeflow.org/last-activity"] = rng.choice([
            _stamp(now - rng.randrange(0, 4000)), _stamp(now - rng.randrange(0, 4000)),
            "2021-08-30T15:37:36.990063Z", "1900-08-30T15:37:36.990063Z", "should-fail",
            "2026-11-06T11:00:00+02:00", ""])
    if rng.random() < 0.75:
        a["notebooks.kubeflow.org/last_activity_check_timestamp"] = rng.choice([
            _stamp(now - rng.randrange(0, 200)), _stamp(now - 60), _stamp(now - 61), "garbage",
            _stamp(now + 30)])
    if rng.random() < 0.5:
        a["other"] = "kept"
    r = rng.random()
    if r < 0.8:
        nb["metadata"]["annotations"] = a
    elif r < 0.9 and not a:
        nb["metadata"]["annotations"] = None

    def acts(n, with_state):
        out = []
        for _ in range(n):
            e = {"id": "k%d" % rng.randrange(1000), "name": "python3",
                 "last_activity": rng.choice([_stamp(now - rng.randrange(0, 5000)),
                                              "2026-11-06T11:59:00.123456Z",
                                              "2026-11-06T13:30:00+02:00", "bad-time"] if rng.random() < 0.15
                                             else [_stamp(now - rng.randrange(0, 5000))])}
            if with_state:
                e["execution_state"] = "idle" if rng.random() < 0.8 else rng.choice(["busy", "starting"])
                e["connections"] = rng.randrange(3)
            out.append(e)
        return out
    kernels = None if rng.random() < 0.15 else acts(rng.randint(0, 3), True)
    terminals = None if rng.random() < 0.3 else acts(rng.randint(0, 2), False)
    return nb, rng.random() > 0.1, kernels, terminals


# ---------------------------------------------------------------- PodDefault admission (C3)


def poddefaults_for_namespace(rng, ns, count):
    """`count` PodDefaults in namespace `ns`: the add-gcp-secret shape of
    admission-webhook/README.md (1 matchLabel, 1 volume, 1 volumeMount), 25 % with
    env vars, 10 % with a sidecar (admission-webhook/examples/
    poddefault-inject-containers.yaml), 5 % conflicting on purpose, a few with
    matchExpressions, tolerations, annotations, labels, command/args."""
    out = []
    for j in range(count):
        name = "pd-%s-%03d" % (ns, j)
        sel = {"matchLabels": {"pd-%d" % j: "true"}}
        r = rng.random()
        if r < 0.08:
            sel = {"matchExpressions": [{"key": "tier", "operator": rng.choice(["In", "NotIn"]),
                                         "values": ["gpu", "cpu"]}]}
        elif r < 0.12:
            sel = {"matchExpressions": [{"key": "pd-%d" % j, "operator": rng.choice(["Exists", "DoesNotExist"])}]}
        elif r < 0.125:
            sel = {}
        spec = {"selector": sel, "desc": "inject %d" % j,
                "volumes": [{"name": "secret-%d" % j, "secret": {"secretName": "user-gcp-sa"}}],
                "volumeMounts": [{"name": "secret-%d" % j, "mountPath": "/secret/%d" % j}]}
        if rng.random() < 0.25:
            spec["env"] = [{"name": "PD_%d_%d" % (j, k), "value": "v%d" % k} for k in range(rng.randint(2, 4))]
        if rng.random() < 0.1:
            spec["sidecars"] = [{"name": "sidecar-%d" % j, "image": "busybox", "command": ["sleep", "3600"]}]
        if rng.random() < 0.05:
            spec["initContainers"] = [{"name": "init-%d" % j, "image": "busybox"}]
        if rng.random() < 0.1:
            spec["tolerations"] = [{"key": "tol-%d" % j, "operator": "Exists", "effect": "NoSchedule"}]
        if rng.random() < 0.15:
            spec["annotations"] = {"pd-ann-%d" % j: "x"}
        if rng.random() < 0.1:
            spec["labels"] = {"pd-label-%d" % j: "y"}
        if rng.random() < 0.08:
            spec["serviceAccountName"] = "sa-%d" % j
        if rng.random() < 0.05:
            spec["automountServiceAccountToken"] = rng.random() < 0.5
        if rng.random() < 0.06:
            spec["command"] = ["/bin/pd-%d" % j]
        if rng.random() < 0.06:
            spec["args"] = ["--from-pd", str(j)]
        if rng.random() < 0.08:
            spec["envFrom"] = [{"configMapRef": {"name": "cm-%d" % j}}]
        if rng.random() < 0.05:
            spec["imagePullSecrets"] = [{"name": "pull-%d" % j}]
        if rng.random() < 0.05:                      # conflicts with what pods / other PodDefaults carry
            spec["env"] = spec.get("env", []) + [{"name": "SHARED", "value": "pd-%d" % j}]
            spec["volumeMounts"].append({"name": "clash-%d" % j, "mountPath": "/home/jovyan"})
        if rng.random() < 0.0015:
            spec["selector"] = {"matchExpressions": [{"key": "a", "operator": "Bogus"}]}
        out.append({"apiVersion": "kubeflow.org/v1alpha1", "kind": "PodDefault",
                    "metadata": {"name": name, "namespace": ns, "resourceVersion": str(5000 + j)},
                    "spec": spec})
    return out


def admission_pod(rng, i, ns, n_pds, match_rate=0.3, per_poddefault=False):
    """Pod of one admission request, shaped like a Notebook StatefulSet's pod
    (1-3 containers, 0-8 env, 0-4 mounts, 3-8 labels).  match_rate: per_poddefault=True reads
    SURVEY.md 8(d)'s "match rate 30 %" literally, every PodDefault of the namespace matches a
    Pod with that probability (30 of 100 on average); False thins it by sqrt(n) so that a Pod
    matches about three at every n (a cluster whose PodDefaults are opt-in features)."""
    labels = {"statefulset": "nb-%06d" % i, "notebook-name": "nb-%06d" % i, "app": "nb"}
    for j in range(n_pds):
        if rng.random() < (match_rate if per_poddefault else match_rate / max(1, n_pds ** 0.5)):
            labels["pd-%d" % j] = "true"
    if rng.random() < 0.3:
        labels["tier"] = rng.choice(["gpu", "cpu", "other"])
    ctrs = []
    for k in range(rng.randint(1, 3)):
        c = {"name": ["nb", "oauth-proxy", "istio-proxy"][k], "image": "img-%d" % k}
        if rng.random() < 0.7:
            c["env"] = [{"name": "E%d" % e, "value": str(e)} for e in range(rng.randint(0, 8))]
            if rng.random() < 0.2:
                c["env"].append({"name": "SHARED", "value": "pod"})
        if rng.random() < 0.7:
            c["volumeMounts"] = [{"name": "vol-%d" % v, "mountPath": ["/home/jovyan", "/dev/shm", "/data", "/x"][v]}
                                 for v in range(rng.randint(0, 4))]
        if rng.random() < 0.2:
            c["command"] = ["start.sh"]
        if rng.random() < 0.1:
            c["args"] = None
        ctrs.append(c)
    md = {"name": "nb-%06d-0" % i, "namespace": ns, "labels": labels}
    r = rng.random()
    if r < 0.5:
        md["annotations"] = {"sidecar.istio.io/inject": "false"}
    if r < 0.04:
        md["annotations"] = {"poddefault.admission.kubeflow.org/exclude": "true"}
    elif r < 0.06:
        md["annotations"] = {"kubernetes.io/config.mirror": "abc"}
    elif r < 0.1:
        md["annotations"] = {"poddefault.admission.kubeflow.org/poddefault-pd-%s-000" % ns: "old-rv"}
    spec = {"containers": ctrs, "serviceAccountName": "default-editor"}
    if rng.random() < 0.6:
        spec["volumes"] = [{"name": "vol-%d" % v, "emptyDir": {}} for v in range(rng.randint(0, 3))]
    if rng.random() < 0.2:
        spec["tolerations"] = [{"key": "node", "operator": "Exists"}]
    return {"apiVersion": "v1", "kind": "Pod", "metadata": md, "spec": spec}


def admission_workload(n_pods, pds_per_ns, seed=SEED, pods_per_ns=20, per_poddefault=False):
    """C3: (pods, poddefaults, ranges) — ranges[i] = [begin, end) of pods[i]'s
    namespace inside the PodDefault list."""
    rng = random.Random(seed + pds_per_ns)
    n_ns = max(1, (n_pods + pods_per_ns - 1) // pods_per_ns)
    pds, ns_range = [], []
    for k in range(n_ns):
        lst = poddefaults_for_namespace(rng, "ns-%03d" % k, pds_per_ns)
        ns_range.append((len(pds), len(pds) + len(lst)))
        pds += lst
    pods, ranges = [], []
    for i in range(n_pods):
        k = i // pods_per_ns
        pods.append(admission_pod(rng, i, "ns-%03d" % k, pds_per_ns, per_poddefault=per_poddefault))
        ranges.append(ns_range[k])
    return pods, pds, ranges


def profile_cr(rng, i):
    """C4: Profile user-%03d owned by user-%03d@example.com; 10 % with a ResourceQuota
    (profile-controller/config/samples; BASELINE.json configs[3])."""
    p = {"apiVersion": "kubeflow.org/v1", "kind": "Profile",
         "metadata": {"name": "user-%03d" % i, "uid": "p-%06d" % i, "resourceVersion": str(i)},
         "spec": {"owner": {"kind": "User", "name": "user-%03d@example.com" % i}}}
    r = rng.random()
    if r < 0.1:
        p["spec"]["resourceQuotaSpec"] = {"hard": {"cpu": "4", "memory": "8Gi"}}
    elif r < 0.15:
        p["spec"]["resourceQuotaSpec"] = {"hard": {}}
    elif r < 0.2:
        p["spec"]["owner"] = {"kind": "User", "apiGroup": "rbac.authorization.k8s.io", "name": "a\"b@x"}
    elif r < 0.23:
        del p["spec"]["owner"]
    elif r < 0.26:
        del p["metadata"]["uid"]
    return p


# ---- tensorboard-controller: Tensorboard CRs (tensorboard-controller/api/v1alpha1/tensorboard_types.go)
TENSORBOARD_PATHS = ["pvc://logs-pvc/run1", "pvc://logs-pvc/", "pvc://logs-pvc", "pvc://", "pvc:///x", "pvc://a/b/c/d",
                     "gs://bucket/logs", "s3://bucket/logs", "/cns/cell/home/logs", "/some/legacy/path", "", "relative",
                     "pvc://na\"me/sub\\dir", "gs://b/\u00e9t\u00e9", "pvc://p/\ttab", "pvc:/notpvc", "PVC://upper/x"]


def tensorboard_cr(rng, i):
    """One Tensorboard CR: every kind of log path the controller tells apart, labels with and
    without an "app" key, sometimes no namespace / null labels / no spec."""
    labels = {}
    if rng.random() < 0.7:
        for k in rng.sample(["app", "team", "access-ml-pipeline", "a/b", "x\"y"], rng.randint(0, 4)):
            labels[k] = rng.choice(["true", "v", "", "tb-%d" % i])
    meta = {"name": "tb-%d" % i, "namespace": rng.choice(["user-ns", "kubeflow-user-example-com", "n%d" % (i % 7)]),
            "uid": "3c1f-%04d" % i}
    if labels or rng.random() < 0.2:
        meta["labels"] = labels
    if rng.random() < 0.1:
        del meta["namespace"]
    if rng.random() < 0.05:
        meta["labels"] = None
    tb = {"apiVersion": "tensorboard.kubeflow.org/v1alpha1", "kind": "Tensorboard", "metadata": meta,
          "spec": {"logspath": rng.choice(TENSORBOARD_PATHS)}}
    if rng.random() < 0.3:
        tb["status"] = {"conditions": [{"deploymentState": "Available"}], "readyReplicas": 1}
    if rng.random() < 0.03:
        tb["spec"] = {}
    return tb


def tensorboard_node(rng):
    """The node-name side input of one Tensorboard: absent, empty, or a JSON string."""
    r = rng.random()
    if r < 0.4:
        return b""
    if r < 0.5:
        return b'""'
    return json.dumps(rng.choice(["node-1", "ip-10-0-0-1.ec2.internal", "n\"q"])).encode()


# ---- pvcviewer-controller: PVCViewer CRs (pvcviewer-controller/api/v1alpha1/pvcviewer_types.go)
def pvcviewer_cr(rng, i):
    """One PVCViewer: with and without networking (zero value: no Service / VirtualService),
    a pod spec of its own or the webhook's default shape, rwoScheduling on and off."""
    name = "viewer-%d" % i
    meta = {"name": name, "namespace": rng.choice(["user-ns", "kubeflow-user-example-com", "n%d" % (i % 5)]),
            "uid": "7a2e-%04d" % i}
    if rng.random() < 0.05:
        del meta["namespace"]
    pod = {"containers": [{"name": "pvcviewer", "image": rng.choice(["filebrowser/filebrowser:latest", "img\"q:1"]),
                           "ports": [{"containerPort": 8080, "protocol": "TCP"}],
                           "env": [{"name": "FB_PORT", "value": "8080"}], "workingDir": "/data",
                           "volumeMounts": [{"name": "viewer-volume", "mountPath": "/data"}]}],
           "volumes": [{"name": "viewer-volume", "persistentVolumeClaim": {"claimName": "pvc-%d" % i}}]}
    if rng.random() < 0.3:
        pod["affinity"] = {"podAffinity": {}}
    if rng.random() < 0.3:
        pod["securityContext"] = {"runAsUser": 1234}
    spec = {"pvc": "pvc-%d" % i, "podSpec": pod, "rwoScheduling": rng.random() < 0.5}
    r = rng.random()
    if r < 0.6:
        spec["networking"] = {"targetPort": rng.choice([8080, "http", 0]), "basePrefix": rng.choice(["/pvcviewer", "", "/x\"y"]),
                              "rewrite": rng.choice(["/", ""]), "timeout": rng.choice(["30s", ""])}
        if rng.random() < 0.3:
            del spec["networking"]["rewrite"]
            del spec["networking"]["timeout"]
    elif r < 0.75:
        spec["networking"] = {"targetPort": 0, "basePrefix": "", "rewrite": ""}     # the zero value, spelled out
    elif r < 0.8:
        spec["networking"] = None
    if rng.random() < 0.05:
        del spec["podSpec"]
    return {"apiVersion": "kubeflow.org/v1alpha1", "kind": "PVCViewer", "metadata": meta, "spec": spec}


def pvcviewer_found(rng, obj, viewer):
    """What a Get returns for an object the controller wrote earlier: sometimes NotFound (None),
    else the object with what an API server and other actors add, sometimes drifted, sometimes
    owned by someone else."""
    if obj is None or rng.random() < 0.35:
        return None
    f = json.loads(json.dumps(obj))
    f["metadata"].update({"resourceVersion": str(rng.randint(1, 99999)), "generation": 2,
                          "creationTimestamp": "2026-11-07T00:00:00Z"})
    if rng.random() < 0.5:
        f["metadata"]["labels"] = dict(f["metadata"].get("labels") or {}, extra="x")
    f["status"] = {"observedGeneration": 2}
    spec = f.setdefault("spec", {})
    if f.get("kind") == "Deployment":
        spec.update({"replicas": rng.choice([0, 1, 2]), "revisionHistoryLimit": 10, "strategy": {"type": "RollingUpdate"}})
        if rng.random() < 0.5:
            spec["template"]["spec"]["affinity"] = {"nodeAffinity": {"requiredDuringSchedulingIgnoredDuringExecution": {"nodeSelectorTerms": []}}}
        elif rng.random() < 0.3:
            spec["template"]["spec"].pop("affinity", None)
    elif f.get("kind") == "Service":
        spec.update({"clusterIP": "10.0.0.%d" % rng.randint(1, 250), "sessionAffinity": "None", "type": "NodePort"})
    r = rng.random()
    refs = f["metadata"].get("ownerReferences") or []
    if r < 0.15:
        f["metadata"].pop("ownerReferences", None)
    elif r < 0.3:
        f["metadata"]["ownerReferences"] = [{"apiVersion": "v1", "kind": "ConfigMap", "name": "other", "uid": "u-other"}] + refs
    elif r < 0.36:
        f["metadata"]["ownerReferences"] = [{"apiVersion": "apps/v1", "kind": "ReplicaSet", "name": "rs", "uid": "u-rs",
                                             "controller": True}]
    elif r < 0.45 and refs:
        refs[0]["blockOwnerDeletion"] = False      # the same owner, to be replaced in place
        f["metadata"]["ownerReferences"] = refs + [{"apiVersion": "v1", "kind": "Secret", "name": "s", "uid": "u-s"}]
    return f
