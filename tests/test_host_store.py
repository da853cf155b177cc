"""CPU: the C++ object store (kubeflow-rm_b200/host/store.cc) behaves like the client calls
of NotebookReconciler.Reconcile expect (NotFound / AlreadyExists / Conflict, resourceVersion,
status subresource)."""
import importlib
import json

import __graft_entry__ as graft


def test_store_semantics(pkg):
    host = importlib.import_module(graft.PKG_NAME + ".host")
    s = host.Store()
    nb = {"apiVersion": "kubeflow.org/v1", "kind": "Notebook", "metadata": {"name": "a", "namespace": "ns"},
          "spec": {"x": 1}}
    assert s.get("Notebook", "ns", "a") is None
    assert s.create(json.dumps(nb).encode()) == 0
    assert s.create(json.dumps(nb).encode()) == 2                      # AlreadyExists
    got = json.loads(s.get("Notebook", "ns", "a"))
    assert got["spec"] == {"x": 1} and got["metadata"]["resourceVersion"] == "1" and got["metadata"]["uid"]
    stale = json.loads(json.dumps(got))
    got["spec"]["x"] = 2
    assert s.update(json.dumps(got).encode()) == 0
    stale["spec"]["x"] = 3
    assert s.update(json.dumps(stale).encode()) == 3                   # Conflict: stale resourceVersion
    del stale["metadata"]["resourceVersion"]
    assert s.update(json.dumps(stale).encode()) == 0                   # no version sent: last write wins
    assert json.loads(s.get("Notebook", "ns", "a"))["spec"]["x"] == 3
    assert s.update(json.dumps(dict(nb, metadata={"name": "zz", "namespace": "ns"})).encode()) == 1
    assert s.create(b"[]") == 4 and s.create(b'{"kind":"X"}') == 4
    assert s.delete("Notebook", "ns", "a") == 0 and s.delete("Notebook", "ns", "a") == 1
    c = s.counters()
    assert c["creates"] == 1 and c["updates"] == 2 and c["conflicts"] == 1 and c["deletes"] == 1
    s.close()


def test_merge_patch_and_json_patch(pkg):
    """RFC 7386 and RFC 6902 application on stored objects (Store::Patch / PatchJson): the first is
    what RemoveReconciliationLock sends (odh-notebook-controller/controllers/
    notebook_controller.go:141-145), the second what the webhooks answer with."""
    import importlib
    import json
    import __graft_entry__ as graft
    host = importlib.import_module(graft.PKG_NAME + ".host")
    st = host.Store()
    nb = {"apiVersion": "kubeflow.org/v1", "kind": "Notebook",
          "metadata": {"name": "a", "namespace": "n", "annotations": {"kubeflow-resource-stopped": "odh-notebook-controller-lock",
                                                                      "keep/this": "x"}},
          "spec": {"template": {"spec": {"containers": [{"name": "a", "image": "i"}], "volumes": [{"name": "v"}]}}}}
    assert st.create(json.dumps(nb).encode()) == 0
    assert st.patch("Notebook", "n", "a", b'{"metadata":{"annotations":{"kubeflow-resource-stopped":null}}}') == 0
    got = json.loads(st.get("Notebook", "n", "a"))
    assert got["metadata"]["annotations"] == {"keep/this": "x"} and got["spec"] == nb["spec"]
    assert st.patch("Notebook", "n", "a", b'{"metadata":{"labels":{"l":"1"}},"spec":{"template":{"spec":{"volumes":null,"nodeName":"x"}}}}') == 0
    got = json.loads(st.get("Notebook", "n", "a"))
    assert got["metadata"]["labels"] == {"l": "1"} and "volumes" not in got["spec"]["template"]["spec"]
    assert got["spec"]["template"]["spec"]["nodeName"] == "x" and got["spec"]["template"]["spec"]["containers"] == nb["spec"]["template"]["spec"]["containers"]
    patch = [{"op": "add", "path": "/metadata/annotations/a~1b", "value": "v"},
             {"op": "add", "path": "/spec/template/spec/containers/-", "value": {"name": "side"}},
             {"op": "replace", "path": "/spec/template/spec/containers/0/image", "value": "j"},
             {"op": "remove", "path": "/metadata/labels"},
             {"op": "add", "path": "/spec/template/spec/containers/0", "value": {"name": "first"}}]
    assert st.patch("Notebook", "n", "a", json.dumps(patch).encode(), json_patch=True) == 0
    got = json.loads(st.get("Notebook", "n", "a"))
    assert got["metadata"]["annotations"] == {"keep/this": "x", "a/b": "v"} and "labels" not in got["metadata"]
    assert [c["name"] for c in got["spec"]["template"]["spec"]["containers"]] == ["first", "a", "side"]
    assert got["spec"]["template"]["spec"]["containers"][1]["image"] == "j"
    assert st.patch("Notebook", "n", "a", b'[{"op":"replace","path":"/nope/x","value":1}]', json_patch=True) == 4
    assert st.patch("Notebook", "n", "missing", b"{}") == 1
    st.close()


def test_workqueue_semantics(pkg):
    """One entry per key, never handed out while processing, RequeueAfter and the per-key back-off."""
    import importlib
    import __graft_entry__ as graft
    host = importlib.import_module(graft.PKG_NAME + ".host")
    q = host.WorkQueue()
    for _ in range(3):
        q.add("ns", "a")
    q.add("ns", "b")
    assert len(q) == 2
    got = q.take(10, 0.0)
    assert got == [("ns", "a"), ("ns", "b")] and len(q) == 0
    q.add("ns", "a")                       # an event for a key that is being reconciled ...
    assert q.take(10, 0.0) == []           # ... is not handed out concurrently ...
    q.done("ns", "a")
    assert q.take(10, 0.0) == [("ns", "a")]   # ... but is not lost either
    q.done("ns", "a")
    q.done("ns", "b")
    q.add_after("ns", "c", 100.0, 60.0)    # ctrl.Result{RequeueAfter: 1 min}
    assert q.take(10, 159.0) == [] and q.next_ready() == 160.0
    assert q.take(10, 160.0) == [("ns", "c")]
    q.done("ns", "c")
    delays = [q.add_rate_limited("ns", "d", 0.0) for _ in range(4)]
    assert delays == [0.005, 0.01, 0.02, 0.04]
    assert q.take(10, 1.0) == [("ns", "d")]     # the four adds collapse into one entry
    q.done("ns", "d")
    q.forget("ns", "d")
    assert q.add_rate_limited("ns", "d", 0.0) == 0.005
    for _ in range(30):
        last = q.add_rate_limited("ns", "e", 0.0)
    assert last == 1000.0
    q.close()


def test_kfam_binding_names(pkg):
    """getBindingName against the reference's table (access-management/kfam/bindings_test.go:25-41;
    the Binding there is User / ClusterRole "edit"), oracle and C++ alike."""
    import importlib
    import random
    import __graft_entry__ as graft
    from oracle import profile as oprofile
    host = importlib.import_module(graft.PKG_NAME + ".host")
    table = [("lalith.vaka@zq.msds.kp.org", "user-lalith-vaka-zq-msds-kp-org-clusterrole-edit"),
             ("397401@zq.msds.kp.org", "user-397401-zq-msds-kp-org-clusterrole-edit"),
             ("lalith.397401@zq.msds.kp.org", "user-lalith-397401-zq-msds-kp-org-clusterrole-edit"),
             ("397401.vaka@zq.msds.kp.org", "user-397401-vaka-zq-msds-kp-org-clusterrole-edit"),
             ("i397401@zq.msds.kp.org", "user-i397401-zq-msds-kp-org-clusterrole-edit")]
    for name, want in table:
        assert oprofile.kfam_binding_name("User", name, "ClusterRole", "edit") == want
        assert host.kfam_binding_name("User", name, "ClusterRole", "edit") == want
    rng = random.Random(1)
    for _ in range(500):
        name = "".join(rng.choice("aZ09.@_-+ é/~") for _ in range(rng.randrange(0, 20)))
        kind = rng.choice(["User", "Group", "ServiceAccount"])
        assert host.kfam_binding_name(kind, name, "ClusterRole", "kubeflow-edit") == \
            oprofile.kfam_binding_name(kind, name, "ClusterRole", "kubeflow-edit")
