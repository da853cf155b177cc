"""GPU: ProfileReconciler (C++ host, kubeflow-rm_b200/host/profile_reconciler.cc) over the
in-memory store, against the flow of profile-controller/controllers/profile_controller.go:105-334
and the oracle's desired objects: everything created on the first pass, nothing but the
Profile's own write on the second (PatchDefaultPluginSpec :735 writes it every time), drift
brought back, a ResourceQuota the Profile no longer wants deleted, a foreign Namespace refused
with a Failed condition, the default plugin patched in, the finalizer added and removed."""
import importlib
import json
import random

import pytest

import __graft_entry__ as graft
from oracle import gojson, profile as oprof

pytestmark = pytest.mark.gpu

LABELS = {"katib.kubeflow.org/metrics-collector-injection": "enabled", "serving.kubeflow.org/inferenceservice": "enabled",
          "app.kubernetes.io/part-of": "kubeflow-profile", "istio-injection": ""}
KINDS = ("Namespace", "AuthorizationPolicy", "ServiceAccount", "RoleBinding", "ServiceAccount", "RoleBinding",
         "RoleBinding", "ResourceQuota")


def stored(store, obj):
    m = obj["metadata"]
    raw = store.get(obj["kind"], m.get("namespace", ""), m["name"])
    return None if raw is None else json.loads(raw)


def strip(o):
    o = json.loads(json.dumps(o))
    for k in ("resourceVersion", "uid", "creationTimestamp", "generation"):
        o["metadata"].pop(k, None)
    return o


def test_profile_reconciler(pkg, wl):
    host = importlib.import_module(graft.PKG_NAME + ".host")
    env = {"NAMESPACE_LABELS_JSON": json.dumps(LABELS)}
    ctx = pkg.Context(0, env=env)
    store = host.Store()
    rec = host.ProfileReconciler(store, ctx, workload_identity="", default_labels=LABELS)
    rng = random.Random(7)
    profiles = [wl.profile_cr(rng, i) for i in range(60)]
    for p in profiles:
        p["metadata"].setdefault("uid", "uid-" + p["metadata"]["name"])
        assert store.create(wl.dumps(p)) == 0
    names = [p["metadata"]["name"] for p in profiles]
    # ---- first pass: every object created, the finalizer added
    rc, st = rec.reconcile(names)
    assert rc == 0 and st["reconciled"] == 60 and st["finalizers_added"] == 60 and st["updates"] == 0
    n_quota = 0
    for p in profiles:
        cur = json.loads(store.get("Profile", "", p["metadata"]["name"]))
        assert cur["metadata"]["finalizers"] == ["profile-finalizer"]
        want = oprof.profile_objects(cur, {}, LABELS)
        for k, w in enumerate(want):
            if w is None:
                assert store.get("ResourceQuota", p["metadata"]["name"], "kf-resource-quota") is None
                continue
            got = stored(store, w)
            assert got is not None and gojson.canon(strip(got)) == gojson.canon(w), (k, got, w)
            n_quota += k == 7
    assert st["creates"] == 60 * 7 + n_quota and 0 < n_quota < 60
    # ---- second pass: nothing to do but the Profile's unconditional write (:735)
    store.counters(reset=True)
    rc, st = rec.reconcile(names + ["no-such-profile"])
    assert rc == 0 and st["not_found"] == 1 and st["creates"] == st["updates"] == st["deletes"] == 0
    assert st["profile_writes"] == 60 and st["finalizers_added"] == 0
    # ---- drift: a RoleBinding's subjects, an AuthorizationPolicy's spec, namespace labels, a quota that is no longer wanted
    a, b, c = names[0], names[1], names[2]
    rb = json.loads(store.get("RoleBinding", a, "namespaceAdmin"))
    rb["subjects"] = [{"kind": "User", "name": "mallory"}]
    rb["metadata"]["labels"] = {"kept": "yes"}
    assert store.update(wl.dumps(rb)) == 0
    ap = json.loads(store.get("AuthorizationPolicy", b, "ns-owner-access-istio"))
    ap["spec"]["rules"] = ap["spec"]["rules"][:1]
    assert store.update(wl.dumps(ap)) == 0
    ns = json.loads(store.get("Namespace", "", c))
    ns["metadata"]["labels"] = {"user-label": "stays", "istio-injection": "enabled"}
    assert store.update(wl.dumps(ns)) == 0
    with_quota = next(p for p in profiles if (p["spec"].get("resourceQuotaSpec") or {}).get("hard"))
    qn = with_quota["metadata"]["name"]
    cur = json.loads(store.get("Profile", "", qn))
    cur["spec"].pop("resourceQuotaSpec")
    assert store.update(wl.dumps(cur)) == 0
    rc, st = rec.reconcile(names)
    assert rc == 0 and st["updates"] == 3 and st["deletes"] == 1 and st["creates"] == 0
    rb2 = json.loads(store.get("RoleBinding", a, "namespaceAdmin"))
    want = oprof.profile_objects(json.loads(store.get("Profile", "", a)), {}, LABELS)[6]
    assert rb2["subjects"] == want["subjects"] and rb2["roleRef"] == want["roleRef"] and rb2["metadata"]["labels"] == {"kept": "yes"}
    ap2 = json.loads(store.get("AuthorizationPolicy", b, "ns-owner-access-istio"))
    assert gojson.canon(ap2["spec"]) == gojson.canon(oprof.profile_objects(json.loads(store.get("Profile", "", b)), {}, LABELS)[1]["spec"])
    ns2 = json.loads(store.get("Namespace", "", c))
    assert ns2["metadata"]["labels"] == oprof.set_namespace_labels({"user-label": "stays", "istio-injection": "enabled"}, LABELS)
    assert store.get("ResourceQuota", qn, "kf-resource-quota") is None
    # ---- a namespace that belongs to someone else is not taken over
    store.create(wl.dumps({"apiVersion": "v1", "kind": "Namespace", "metadata": {"name": "taken", "annotations": {"owner": "someone@else"}}}))
    store.create(wl.dumps({"apiVersion": "v1", "kind": "Namespace", "metadata": {"name": "unlabelled"}}))
    for n_ in ("taken", "unlabelled"):
        store.create(wl.dumps({"apiVersion": "kubeflow.org/v1", "kind": "Profile", "metadata": {"name": n_, "uid": "u-" + n_},
                               "spec": {"owner": {"kind": "User", "name": "alice@example.com"}}}))
    rc, st = rec.reconcile(["taken", "unlabelled"])
    assert rc == 0 and st["rejected"] == 2 and st["creates"] == 0 and st["finalizers_added"] == 0
    for n_ in ("taken", "unlabelled"):
        cur = json.loads(store.get("Profile", "", n_))
        assert cur["status"]["conditions"] == [{"type": "Failed",
                                                "message": "namespace already exist, but not owned by profile creator alice@example.com"}]
        assert store.get("ServiceAccount", n_, "default-editor") is None
    # ---- the default plugin and the plugins' turn; then deletion: revoke, finalizer gone
    rec2 = host.ProfileReconciler(store, ctx, workload_identity="sa@project.iam.gserviceaccount.com", default_labels=LABELS)
    rc, st = rec2.reconcile(names[:5])
    assert rc == 0 and st["plugins_applied"] >= 5
    for n_ in names[:5]:
        cur = json.loads(store.get("Profile", "", n_))
        wi = [p for p in cur["spec"]["plugins"] if p["kind"] == "WorkloadIdentity"]
        assert len(wi) == 1
    first = json.loads(store.get("Profile", "", names[0]))
    if not any(p.get("spec", {}).get("gcpServiceAccount") == "sa@project.iam.gserviceaccount.com" for p in first["spec"]["plugins"]):
        assert any(p["kind"] == "WorkloadIdentity" for p in profiles[0]["spec"].get("plugins", []))   # the user's own one wins
    rc, st2 = rec2.reconcile(names[:5])          # a second time: no second copy
    assert all(len([p for p in json.loads(store.get("Profile", "", n_))["spec"]["plugins"] if p["kind"] == "WorkloadIdentity"]) == 1
               for n_ in names[:5])
    cur = json.loads(store.get("Profile", "", names[3]))
    cur["metadata"]["deletionTimestamp"] = "2026-11-07T12:00:00Z"
    cur["metadata"]["finalizers"] = ["other-finalizer", "profile-finalizer"]
    assert store.update(wl.dumps(cur)) == 0
    rc, st = rec2.reconcile([names[3]])
    assert rc == 0 and st["finalizers_removed"] == 1 and st["plugins_revoked"] >= 1
    assert json.loads(store.get("Profile", "", names[3]))["metadata"]["finalizers"] == ["other-finalizer"]
    rec.close()
    rec2.close()
    store.close()
    ctx.close()
