"""GPU: the C++ NotebookReconciler over the in-memory store, read like the reference's own
envtest spec (notebook_controller_bdd_test.go:32-89: create a Notebook, a StatefulSet appears)
but asserting the content too, against the oracle."""
import importlib
import json

import pytest

import __graft_entry__ as graft
from oracle import gojson, nbc

pytestmark = pytest.mark.gpu
NOW = 1793966400
ENV = {"USE_ISTIO": "true"}


def _strip(o):
    o = json.loads(json.dumps(o))
    for k in ("resourceVersion", "uid"):
        o["metadata"].pop(k, None)
    return o


def test_reconcile_like_envtest(pkg, wl):
    host = importlib.import_module(graft.PKG_NAME + ".host")
    ctx = pkg.Context(0, env=ENV)
    store = host.Store()
    rec = host.NotebookReconciler(store, ctx, use_istio=True)
    nbs = [wl.notebook_c1(0)] + wl.corpus("c2", 200, start=1) + wl.corpus("loadtest", 50, start=300)
    for nb in nbs:
        assert store.create(wl.dumps(nb)) == 0
    keys = [(nb["metadata"]["namespace"], nb["metadata"]["name"]) for nb in nbs]
    keys.append(("default", "does-not-exist"))
    rc, st = rec.reconcile(keys, NOW)
    assert rc == 0 and st["reconciled"] == len(nbs) and st["not_found"] == 1
    assert st["creates"] == 3 * len(nbs) and st["updates"] == 0 and st["status_updates"] == len(nbs)
    # "By checking that the Notebook has statefulset" -- and that it is the right one
    for nb in nbs:
        stored = json.loads(store.get("Notebook", nb["metadata"]["namespace"], nb["metadata"]["name"]))
        ns, name = nb["metadata"]["namespace"], nb["metadata"]["name"]
        exp = [nbc.generate_statefulset(stored, ENV), nbc.generate_service(stored, ENV),
               nbc.generate_virtualservice(stored, ENV)]
        got = [store.get("StatefulSet", ns, name), store.get("Service", ns, name),
               store.get("VirtualService", ns, "notebook-%s-%s" % (ns, name))]
        for e, g in zip(exp, got):
            nbc.set_controller_reference(stored, e)
            assert g is not None
            assert gojson.canon(_strip(json.loads(g))) == gojson.canon(e)
        assert stored["status"] == {"conditions": [], "readyReplicas": 0, "containerState": {}}
    # level-triggered and idempotent: a second pass writes only the unconditional status (:287)
    store.counters(reset=True)
    rc, st = rec.reconcile(keys, NOW)
    assert rc == 0 and st["creates"] == 0 and st["updates"] == 0 and st["status_updates"] == len(nbs)
    # the culler stops a notebook: replicas go to 0 on the next reconcile (:410-412)
    ns, name = keys[1]
    nb = json.loads(store.get("Notebook", ns, name))
    nb["metadata"].setdefault("annotations", {})["kubeflow-resource-stopped"] = "2026-11-06T12:00:00Z"
    assert store.update(json.dumps(nb).encode()) == 0
    # someone deletes an owned Service: it comes back
    assert store.delete("Service", *keys[2]) == 0
    # a terminating Notebook is left alone (:135-137)
    nb3 = json.loads(store.get("Notebook", *keys[3]))
    nb3["metadata"]["deletionTimestamp"] = "2026-11-06T12:00:00Z"
    assert store.update(json.dumps(nb3).encode()) == 0
    assert store.delete("StatefulSet", *keys[3]) == 0
    rc, st = rec.reconcile(keys, NOW)
    assert rc == 0 and st["updates"] == 1 and st["creates"] == 1 and st["terminating"] == 1
    assert json.loads(store.get("StatefulSet", ns, name))["spec"]["replicas"] == 0
    assert store.get("Service", *keys[2]) is not None and store.get("StatefulSet", *keys[3]) is None
    # write-if-changed status: the steady state becomes write-free
    rec2 = host.NotebookReconciler(store, ctx, use_istio=True, status_write_if_changed=True)
    store.counters(reset=True)
    rc, st = rec2.reconcile(keys, NOW)
    c = store.counters()
    assert rc == 0 and st["status_updates"] == 0 and c["creates"] + c["updates"] + c["status_updates"] == 0
    rec.close(); rec2.close(); store.close(); ctx.close()


def test_pod_restart_and_status_mirroring(pkg, wl):
    host = importlib.import_module(graft.PKG_NAME + ".host")
    ctx = pkg.Context(0, env=ENV)
    store = host.Store()
    rec = host.NotebookReconciler(store, ctx, use_istio=False)
    nb = wl.notebook_c2(5)
    nb["metadata"]["annotations"]["notebooks.opendatahub.io/notebook-restart"] = "true"
    ns, name = nb["metadata"]["namespace"], nb["metadata"]["name"]
    store.create(wl.dumps(nb))
    pod = {"apiVersion": "v1", "kind": "Pod", "metadata": {"name": name + "-0", "namespace": ns},
           "status": {"phase": "Running", "conditions": [{"type": "Ready", "status": "True",
                                                          "lastTransitionTime": "2026-11-06T11:00:00Z"}],
                      "containerStatuses": [{"name": name, "state": {"running": {"startedAt": "2026-11-06T11:00:00Z"}}}]}}
    store.create(wl.dumps(pod))
    rc, st = rec.reconcile([(ns, name)], NOW)
    assert rc == 0 and st["creates"] == 2 and st["pod_restarts"] == 1       # no VirtualService without Istio
    assert store.get("Pod", ns, name + "-0") is None
    stored = json.loads(store.get("Notebook", ns, name))
    assert "notebooks.opendatahub.io/notebook-restart" not in stored["metadata"]["annotations"]
    exp = nbc.create_notebook_status(nb, {}, pod, "2026-11-06T12:00:00Z")
    assert gojson.canon(stored["status"]) == gojson.canon(exp)
    assert store.get("VirtualService", ns, "notebook-%s-%s" % (ns, name)) is None
    # restart asked again with the pod gone: the reference's Delete fails and Reconcile returns the error
    stored["metadata"]["annotations"]["notebooks.opendatahub.io/notebook-restart"] = "true"
    store.update(json.dumps(stored).encode())
    rc, st = rec.reconcile([(ns, name)], NOW)
    assert rc == 1 and st["errors"] == 1
    rec.close(); store.close(); ctx.close()


def test_event_reissue(pkg, wl):
    """Reconcile on an Event key re-issues it on the Notebook (notebook_controller.go:95-119) with
    nbNameFromInvolvedObject as pinned by notebook_controller_test.go:22-92."""
    import os
    host = importlib.import_module(graft.PKG_NAME + ".host")
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "nbc_involved_object.json")))
    ctx = pkg.Context(0, env=ENV)
    store = host.Store()
    rec = host.NotebookReconciler(store, ctx, use_istio=False)
    ns = "test-namespace"
    nb = wl.notebook_c1(1)
    nb["metadata"].update(name="test-notebook", namespace=ns)
    store.create(wl.dumps(nb))
    for p in g["pods"]:
        store.create(wl.dumps(dict(p, kind="Pod", apiVersion="v1")))
    keys = []
    for i, c in enumerate(g["cases"]):
        ev = {"apiVersion": "v1", "kind": "Event", "metadata": {"name": "ev-%d" % i, "namespace": ns},
              "involvedObject": c["involvedObject"], "type": "Warning", "reason": "FailedScheduling",
              "message": "0/1 nodes are available"}
        store.create(wl.dumps(ev))
        keys.append((ns, "ev-%d" % i))
    store.create(wl.dumps({"apiVersion": "v1", "kind": "Event", "metadata": {"name": "ev-x", "namespace": ns},
                           "involvedObject": {"kind": "Node", "name": "n1"}, "type": "Normal", "reason": "r",
                           "message": "m"}))
    # a message with quotes, a backslash and a newline: the re-issued Event is built from the
    # stored text, escapes and all, and must stay valid JSON with the same decoded message
    nasty = 'Back-off pulling image "quay.io/x:1" at C:\\dir\nretrying, said "no"'
    store.create(wl.dumps({"apiVersion": "v1", "kind": "Event", "metadata": {"name": "ev-q", "namespace": ns},
                           "involvedObject": {"kind": "StatefulSet", "name": "test-notebook", "namespace": ns},
                           "type": "Warning", "reason": 'Back"Off', "message": nasty}))
    rc, st = rec.reconcile(keys + [(ns, "ev-x"), (ns, "ev-q")], NOW)
    assert st["events_reissued"] == len(g["cases"]) + 1 and rc == 1 and st["reconciled"] == 0
    out = json.loads(store.get("Event", ns, "test-notebook.ev-q"))
    assert out["message"] == "Reissued from statefulset/test-notebook: " + nasty and out["reason"] == 'Back"Off'
    for i, c in enumerate(g["cases"]):
        out = json.loads(store.get("Event", ns, "%s.ev-%d" % (c["expected"], i)))
        assert out["involvedObject"]["name"] == c["expected"] and out["involvedObject"]["kind"] == "Notebook"
        kind = c["involvedObject"]["kind"].lower()
        assert out["message"] == "Reissued from %s/%s: 0/1 nodes are available" % (kind, c["involvedObject"]["name"])
    rec.close(); store.close(); ctx.close()


def test_rejected_notebooks_stop_before_the_status_write(pkg, wl):
    """`containers: []` and `ports: []`: the reference panics on Containers[0] / Ports[0]
    (notebook_controller.go:455,491) and never reaches Status().Update.  ReconcileBatch
    reports an error for those keys, writes nothing for them, and still reconciles the
    healthy Notebook of the same batch."""
    host = importlib.import_module(graft.PKG_NAME + ".host")
    ctx = pkg.Context(0, env=ENV)
    store = host.Store()
    rec = host.NotebookReconciler(store, ctx, use_istio=True)
    ns = "default"
    good = wl.notebook_c1(1)
    good["metadata"].update(name="good", namespace=ns)
    empty = wl.notebook_c1(2)
    empty["metadata"].update(name="no-containers", namespace=ns)
    empty["spec"]["template"]["spec"]["containers"] = []
    ports = wl.notebook_c1(3)
    ports["metadata"].update(name="no-ports", namespace=ns)
    ports["spec"]["template"]["spec"]["containers"][0]["ports"] = []
    for nb in (good, empty, ports):
        store.create(wl.dumps(nb))
    rc, st = rec.reconcile([(ns, "no-containers"), (ns, "good"), (ns, "no-ports")], NOW)
    assert rc == 2 and st["errors"] == 2 and st["reconciled"] == 3
    assert st["status_updates"] == 1 and st["creates"] == 3
    assert store.get("StatefulSet", ns, "good") is not None
    for name in ("no-containers", "no-ports"):
        assert store.get("StatefulSet", ns, name) is None and store.get("Service", ns, name) is None
        assert "status" not in json.loads(store.get("Notebook", ns, name))
    rec.close(); store.close(); ctx.close()
