"""GPU: the hand-off between the two controllers on the in-memory store, driven
through a work queue: the ODH webhook locks a new Notebook (kubeflow-resource-stopped), so
notebook-controller's first reconcile creates the StatefulSet with replicas 0; the ODH reconciler
creates its objects and removes the lock with a merge patch (odh-notebook-controller/controllers/
notebook_controller.go:118-146, :235-243); the Notebook's update event brings it back to
notebook-controller, which scales to 1 (notebook-controller/controllers/notebook_controller.go:
410-412, :165-176).  Every stored object is checked against the oracle."""
import importlib
import json

import pytest

import __graft_entry__ as graft
from oracle import gojson, nbc, odh, passes

pytestmark = pytest.mark.gpu

ENV = {"USE_ISTIO": "true", "CONTROLLER_NAMESPACE": passes.CTRL_NS}
NOW = 1793966400


def test_lock_replicas0_unlock_replicas1(pkg, wl):
    host = importlib.import_module(graft.PKG_NAME + ".host")
    ctx = pkg.Context(0, env=ENV)
    store = host.Store()
    nbc_rec = host.NotebookReconciler(store, ctx, use_istio=True)
    odh_rec = host.OpenshiftNotebookReconciler(store, ctx, rbac=False)
    q_nbc, q_odh = host.WorkQueue(), host.WorkQueue()
    n = 40
    submitted = wl.corpus("c2", n)
    # admission: the webhook's patch, as the API server applies it before storing
    reviews = [{"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview",
                "request": {"uid": "u%d" % i, "operation": "CREATE", "object": nb,
                            "resource": {"group": "kubeflow.org", "version": "v1", "resource": "notebooks"}}}
               for i, nb in enumerate(submitted)]
    res = ctx.notebook_webhook_review(pkg.Docs([wl.dumps(r) for r in reviews]))
    keys = []
    for nb, raw in zip(submitted, res.docs()):
        import base64
        resp = json.loads(raw)["response"]
        assert resp["allowed"]
        assert store.create(wl.dumps(nb)) == 0
        md = nb["metadata"]
        assert store.patch("Notebook", md["namespace"], md["name"], base64.b64decode(resp["patch"]), json_patch=True) == 0
        stored = json.loads(store.get("Notebook", md["namespace"], md["name"]))
        assert stored["metadata"]["annotations"]["kubeflow-resource-stopped"] == "odh-notebook-controller-lock"
        keys.append((md["namespace"], md["name"]))
        q_nbc.add(*keys[-1])
        q_odh.add(*keys[-1])

    def drain(q, rec, *args):
        batch = q.take(1000, 0.0)
        out = rec.reconcile(batch, *args) if batch else None
        for k in batch:
            q.done(*k)
        return batch, out

    # 1. notebook-controller sees the locked Notebooks: StatefulSets with replicas 0
    batch, (rc, st) = drain(q_nbc, nbc_rec, NOW)
    assert len(batch) == n and rc == 0 and st["creates"] == 3 * n
    for ns, name in keys:
        assert json.loads(store.get("StatefulSet", ns, name))["spec"]["replicas"] == 0
    # 2. the ODH reconciler: its six objects per Notebook, then the lock goes
    batch, (rc, st, slept) = drain(q_odh, odh_rec)
    assert rc == 0 and st["creates"] == 6 * n and st["locks_removed"] == n
    assert st["pull_secret_waits"] == 2 * n and slept == 6.0 * n     # 1 s + 5 s each: no pull secret ever comes
    for (ns, name), nb in zip(keys, submitted):
        stored = json.loads(store.get("Notebook", ns, name))
        assert "kubeflow-resource-stopped" not in stored["metadata"]["annotations"]
        mutated, _ = odh.mutate_notebook(nb, "CREATE")
        want = odh.reconcile_objects(mutated, passes.CTRL_NS, rbac=False, cookie_secret="x")
        kinds = [("NetworkPolicy", name + "-ctrl-np"), ("NetworkPolicy", name + "-oauth-np"), ("Route", name),
                 ("ServiceAccount", name), ("Service", name + "-tls"), ("Secret", name + "-oauth-config")]
        for (kind, oname), w in zip(kinds, want):
            got = json.loads(store.get(kind, ns, oname))
            for k in ("resourceVersion", "uid"):
                got["metadata"].pop(k, None)
            if kind == "Secret":
                assert len(got["stringData"].pop("cookie_secret")) == 32       # the one pinned property
                w["stringData"].pop("cookie_secret")
            assert gojson.canon(got) == gojson.canon(w), (kind, oname)
        q_nbc.add(ns, name)            # the Notebook changed: its update event
    # 3. notebook-controller again: replicas 1, by Update (the replica counts differ)
    batch, (rc, st) = drain(q_nbc, nbc_rec, NOW)
    assert len(batch) == n and rc == 0 and st["updates"] == n and st["creates"] == 0
    for ns, name in keys:
        assert json.loads(store.get("StatefulSet", ns, name))["spec"]["replicas"] == 1
    # 4. steady state: the ODH reconciler finds everything in place
    for k in keys:
        q_odh.add(*k)
    batch, (rc, st, slept) = drain(q_odh, odh_rec)
    assert rc == 0 and st["creates"] == 0 and st["updates"] == 0 and st["locks_removed"] == 0 and slept == 0.0
    # 5. a Route someone edited is brought back, through a conflict
    ns, name = keys[0]
    r = json.loads(store.get("Route", ns, name))
    r["spec"]["to"]["name"] = "hijacked"
    assert store.update(wl.dumps(r)) == 0
    q_odh.add(ns, name)
    batch, (rc, st, slept) = drain(q_odh, odh_rec)
    assert rc == 0 and st["updates"] == 1
    assert json.loads(store.get("Route", ns, name))["spec"]["to"]["name"] == name + "-tls"
    for x in (nbc_rec, odh_rec, q_nbc, q_odh, store, ctx):
        x.close()
