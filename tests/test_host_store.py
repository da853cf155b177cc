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
