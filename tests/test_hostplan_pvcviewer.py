"""CPU: what a PVCViewer reconcile writes (GenPvcViewer, csrc/nbrm_pvcv.cuh) against the oracle's
restatement of reconcileDeployment / reconcileService / reconcileVirtualService
(pvcviewer-controller/controllers/pvcviewer_controller.go:149-336), Create and Update bodies alike;
and the oracle against the assertions of the reference's own specs
(controllers/pvcviewer_controller_test.go, transcribed below with their line numbers)."""
import json
import random

from oracle import gojson, pvcviewer as opv

ENVS = [{}, {"ISTIO_GATEWAY": "istio-system/gw"}, {"ISTIO_GATEWAY": ""}]


def make(wl, n, seed):
    rng = random.Random(seed)
    viewers = [wl.pvcviewer_cr(rng, i) for i in range(n)]
    found, nodes = [], []
    for v in viewers:
        base = opv.viewer_objects(v, {}, nodename="")
        found += [wl.pvcviewer_found(rng, o, v) for o in base]
        nodes.append(rng.choice([b"", b'""', b'"node-7"', b'"n\\"q"']))
    return viewers, found, nodes


def check(viewers, found, nodes, outs, st, env):
    n_create = n_update = n_conflict = 0
    for i, v in enumerate(viewers):
        f = found[3 * i:3 * i + 3] if found is not None else [None] * 3
        node = json.loads(nodes[i]) if nodes is not None and nodes[i] else ""
        try:
            want = opv.viewer_objects(v, env, f, node)
        except opv.AlreadyOwned:
            assert st[i] == 7, (st[i], f)
            assert outs[3 * i:3 * i + 3] == [b"", b"", b""]
            n_conflict += 1
            continue
        assert st[i] == 0, (st[i], v, f)
        for k in range(3):
            if want[k] is None:
                assert outs[3 * i + k] == b""
                continue
            got = json.loads(outs[3 * i + k])
            assert gojson.canon(got) == gojson.canon(want[k]), (k, outs[3 * i + k], want[k])
            if f[k] is None:
                n_create += 1
            else:
                n_update += 1
    return n_create, n_update, n_conflict


def test_against_oracle(hostplan, wl):
    viewers, found, nodes = make(wl, 1500, 31)
    rng = random.Random(32)
    docs = [wl.dumps(v, rng) for v in viewers]
    fdocs = [b"" if f is None else wl.dumps(f, rng) for f in found]
    tot = [0, 0, 0]
    for env in ENVS:
        outs, st = hostplan.pvcviewer(docs, fdocs, nodes, env)
        r = check(viewers, found, nodes, outs, st, env)
        tot = [a + b for a, b in zip(tot, r)]
        outs, st = hostplan.pvcviewer(docs, None, None, env)
        check(viewers, None, None, outs, st, env)
    assert tot[0] > 2000 and tot[1] > 2000 and tot[2] > 100, tot


def test_reference_specs_on_the_oracle():
    """controllers/pvcviewer_controller_test.go, the assertions that concern the objects."""
    viewer = {"apiVersion": "kubeflow.org/v1alpha1", "kind": "PVCViewer",
              "metadata": {"name": "test-pvcviewer", "namespace": "default", "uid": "u1"},
              "spec": {"pvc": "test-pvc", "networking": {"targetPort": 8080, "basePrefix": "/pvcviewer"}}}
    # :70-81 the defaulting webhook creates a pod spec with one container and the PVC's volume
    d = opv.default(viewer)
    assert len(d["spec"]["podSpec"]["containers"]) == 1 and d["spec"]["podSpec"]["containers"][0]["image"]
    assert len(d["spec"]["podSpec"]["volumes"]) == 1
    assert d["spec"]["podSpec"]["volumes"][0]["persistentVolumeClaim"]["claimName"] == "test-pvc"
    # :83-97 defaults from the file named by the environment variable
    d2 = opv.default(viewer, {"containers": [{"name": "test"}], "securityContext": {"runAsUser": 1234}})
    assert d2["spec"]["podSpec"]["containers"][0]["name"] == "test"
    assert d2["spec"]["podSpec"]["securityContext"]["runAsUser"] == 1234
    # :99-112, :114-136 the validating webhook's two messages
    assert opv.validate({"spec": {"pvc": "", "podSpec": d["spec"]["podSpec"]}}) == "PVC name must be specified"
    bad = json.loads(json.dumps(d))
    bad["spec"]["podSpec"]["volumes"][0]["persistentVolumeClaim"]["claimName"] = "other"
    assert opv.validate(bad) == "PVC test-pvc must be used in the podSpec" and opv.validate(d) is None
    # :142-160 Deployment: the viewer's container image, the three labels
    dep, svc, vs = opv.viewer_objects(d, {})
    assert dep["spec"]["template"]["spec"]["containers"][0]["image"] == d["spec"]["podSpec"]["containers"][0]["image"]
    labels = {"app.kubernetes.io/name": "test-pvcviewer", "app.kubernetes.io/instance": "pvcviewer-test-pvcviewer",
              "app.kubernetes.io/part-of": "pvc-viewer"}
    assert dep["metadata"]["labels"] == labels and dep["metadata"]["name"] == "pvcviewer-test-pvcviewer"
    # :191-247 Service: one port "http", 80 -> targetPort
    assert svc["spec"]["ports"] == [{"name": "http", "port": 80, "targetPort": 8080}] and svc["metadata"]["labels"] == labels
    # :249-293 VirtualService: labels, timeout nil, prefix == rewrite == basePrefix/namespace/name/
    http0 = vs["spec"]["http"][0]
    assert vs["metadata"]["labels"] == labels and http0["timeout"] is None
    assert http0["match"][0]["uri"]["prefix"] == "/pvcviewer/default/test-pvcviewer/" == http0["rewrite"]["uri"]
    assert len(vs["metadata"]["ownerReferences"]) == 1
    # :295-334 after an update of rewrite and timeout; still one ownerReference
    d3 = json.loads(json.dumps(d))
    d3["spec"]["networking"].update({"rewrite": "/new", "timeout": "30s"})
    vs2 = opv.reconcile_virtualservice(d3, {}, vs)
    assert vs2["spec"]["http"][0]["timeout"] == "30s" and vs2["spec"]["http"][0]["rewrite"]["uri"] == "/new"
    assert vs2["spec"]["http"][0]["match"][0]["uri"]["prefix"] == "/pvcviewer/default/test-pvcviewer/"
    assert len(vs2["metadata"]["ownerReferences"]) == 1
    # :162-189 an updated pod spec reaches the Deployment
    d3["spec"]["podSpec"]["containers"][0]["image"] = "new-image"
    d3["spec"]["podSpec"]["containers"].append({"name": "second", "image": "new-image"})
    dep2 = opv.reconcile_deployment(d3, dep)
    assert [c["image"] for c in dep2["spec"]["template"]["spec"]["containers"]] == ["new-image", "new-image"]
    # :336-368 status: ready once readyReplicas == replicas, url from the networking
    st = opv.reconcile_status(d, dict(dep, spec=dict(dep["spec"], replicas=1), status={"readyReplicas": 1}))
    assert st["ready"] is True and st["url"] == "/pvcviewer/default/test-pvcviewer/"
    # :371-388 no affinity without a node; :390-436 an affinity for the node of the RWO volume
    rwo = json.loads(json.dumps(d))
    rwo["spec"]["rwoScheduling"] = True
    assert "affinity" not in opv.reconcile_deployment(rwo)["spec"]["template"]["spec"]
    aff = opv.reconcile_deployment(rwo, None, "node-1")["spec"]["template"]["spec"]["affinity"]
    assert aff["nodeAffinity"]["preferredDuringSchedulingIgnoredDuringExecution"][0]["preference"]["matchExpressions"][0]["values"] == ["node-1"]


DEFAULT_FILE = {"containers": [{"name": "test", "image": "img"}], "securityContext": {"runAsUser": 1234},
                "volumes": [{"name": "scratch", "emptyDir": {}}]}


def webhook_viewers(wl, n, seed):
    rng = random.Random(seed)
    out = []
    for i in range(n):
        v = wl.pvcviewer_cr(rng, i)
        r = rng.random()
        if "podSpec" not in v["spec"]:
            pass
        elif r < 0.35:
            v["spec"].pop("podSpec", None)
        elif r < 0.45:
            v["spec"]["podSpec"] = rng.choice([{}, None, {"containers": None, "volumes": None}])
        elif r < 0.55:
            v["spec"]["podSpec"]["volumes"][0]["persistentVolumeClaim"]["claimName"] = "other"
        elif r < 0.6:
            v["spec"]["podSpec"]["volumes"] = rng.choice([[], None, [{"name": "x", "emptyDir": {}}]])
        if rng.random() < 0.1:
            v["spec"]["pvc"] = rng.choice(["", None])
        if rng.random() < 0.03:
            del v["spec"]
        out.append(v)
    return out


def check_webhook(viewers, outs, st, res, default_file):
    codes = {None: 0, "PVC name must be specified": 1, "PodSpec must be specified": 2}
    seen = [0, 0, 0, 0]
    for v, o, s, r in zip(viewers, outs, st, res):
        assert s == 0
        want = opv.default(v, default_file)
        assert gojson.canon(json.loads(o)) == gojson.canon(want), (o, want)
        msg = opv.validate(want)
        code = codes.get(msg, 3)
        assert r == code, (r, msg, v)
        seen[code] += 1
    return seen


def test_webhook_default_and_validate(hostplan, wl):
    viewers = webhook_viewers(wl, 1500, 51)
    rng = random.Random(52)
    docs = [wl.dumps(v, rng) for v in viewers]
    outs, st, res = hostplan.pvcviewer_webhook(docs, {})
    seen = check_webhook(viewers, outs, st, res, None)
    assert seen[0] > 500 and seen[1] > 50 and seen[3] > 50, seen
    for f in (DEFAULT_FILE, {"containers": None}, {"containers": [{"name": "only"}]}):
        outs, st, res = hostplan.pvcviewer_webhook(docs, {"PVCVIEWER_DEFAULT_PODSPEC_JSON": json.dumps(f)})
        check_webhook(viewers, outs, st, res, f)


def status_cases(wl, n, seed):
    rng = random.Random(seed)
    viewers, deps = [], []
    cond_a = {"type": "Available", "status": "True", "lastUpdateTime": "2026-11-07T10:00:00Z"}
    cond_b = {"type": "Progressing", "status": "True", "reason": "NewReplicaSetAvailable"}
    for i in range(n):
        v = wl.pvcviewer_cr(rng, i)
        v["status"] = rng.choice([None, {}, {"ready": False}, {"ready": True, "conditions": [cond_a], "url": "/old/"},
                                  {"conditions": [cond_b, cond_a]}])
        if v["status"] is None:
            del v["status"]
        r = rng.random()
        if r < 0.2:
            deps.append(None)
        else:
            d = {"apiVersion": "apps/v1", "kind": "Deployment", "metadata": {"name": "pvcviewer-" + v["metadata"]["name"]},
                 "spec": {"replicas": rng.choice([0, 1, 2])},
                 "status": {"readyReplicas": rng.choice([0, 1, 2]), "conditions": rng.choice([[], [cond_a], [cond_b, cond_a], [dict(cond_a, status="False")]])}}
            if rng.random() < 0.2:
                del d["status"]["readyReplicas"]
            if rng.random() < 0.1:
                d["status"] = {}
            deps.append(d)
        viewers.append(v)
    return viewers, deps


def check_status(viewers, deps, outs, st):
    for v, d, o, s in zip(viewers, deps, outs, st):
        assert s == 0
        assert gojson.canon(json.loads(o)) == gojson.canon(opv.reconcile_status(v, d)), (o, v, d)


def test_reconcile_status(hostplan, wl):
    viewers, deps = status_cases(wl, 1500, 81)
    rng = random.Random(82)
    outs, st = hostplan.pvcviewer_status([wl.dumps(v, rng) for v in viewers],
                                         [b"" if d is None else wl.dumps(d, rng) for d in deps])
    check_status(viewers, deps, outs, st)
    # a Deployment without spec.replicas: the reference dereferences nil
    outs, st = hostplan.pvcviewer_status([wl.dumps(viewers[0])], [b'{"kind":"Deployment","spec":{}}'])
    assert list(st) == [5] and outs == [b""]
