"""CPU: the tensorboard-controller's desired objects (GenTensorboard, csrc/nbrm_tb.cuh) against the
oracle's restatement of generateDeployment / generateService / generateVirtualService
(tensorboard-controller/controllers/tensorboard_controller.go:167-378) and of its path helpers."""
import json
import random

import numpy as np

from oracle import gojson, tensorboard as otb

ENV = {"TENSORBOARD_IMAGE": "tensorflow/tensorflow:2.5.1", "ISTIO_GATEWAY": "kubeflow/kubeflow-gateway",
       "ISTIO_HOST": "*"}
ENVS = [ENV,
        dict(ENV, RWO_PVC_SCHEDULING="true"),
        dict(ENV, RWO_PVC_SCHEDULING="False", ISTIO_HOST="tb.example.com", TENSORBOARD_IMAGE="reg/\"tb\":1"),
        dict(ENV, RWO_PVC_SCHEDULING="TRUE", TENSORBOARD_IMAGE="", ISTIO_GATEWAY="")]

def check(tbs, nodes, outs, st, env, owner):
    n_ok = 0
    for i, tb in enumerate(tbs):
        node = json.loads(nodes[i]) if nodes is not None and nodes[i] else ""
        try:
            want = otb.tensorboard_objects(tb, env, node, owner)
        except otb.ReferenceError_:
            assert st[i] == 10, (st[i], tb, env)
            assert outs[3 * i:3 * i + 3] == [b"", b"", b""]
            continue
        assert st[i] == 0, (st[i], tb)
        for k in range(3):
            got = json.loads(outs[3 * i + k])
            assert gojson.canon(got) == gojson.canon(want[k]), (k, outs[3 * i + k], want[k])
        n_ok += 1
    return n_ok


def test_path_helpers():
    assert otb.extract_pvc_name("pvc://a/b/c") == "a" and otb.extract_pvc_subpath("pvc://a/b/c") == "b/c"
    assert otb.extract_pvc_name("pvc://a") == "a" and otb.extract_pvc_subpath("pvc://a") == ""
    assert otb.extract_pvc_subpath("pvc://a/") == "" and otb.extract_pvc_name("pvc:///x") == ""
    assert otb.is_cloud_path("/cns/x") and otb.is_cloud_path("s3://b") and not otb.is_cloud_path("pvc://x")


def test_against_oracle(hostplan, wl):
    rng = random.Random(11)
    tbs = [wl.tensorboard_cr(rng, i) for i in range(1200)]
    docs = [wl.dumps(t, rng) for t in tbs]
    nodes = [wl.tensorboard_node(rng) for _ in tbs]
    seen = 0
    for env in ENVS:
        for flags in (0, 8):
            outs, st = hostplan.tensorboard(docs, nodes, env, flags=flags)
            seen += check(tbs, nodes, outs, st, env, bool(flags))
            outs2, st2 = hostplan.tensorboard(docs, None, env, flags=flags)
            check(tbs, None, outs2, st2, env, bool(flags))
    assert seen > 8000


def test_missing_environment(hostplan, wl):
    rng = random.Random(12)
    tbs = [wl.tensorboard_cr(rng, i) for i in range(60)]
    docs = [wl.dumps(t) for t in tbs]
    for missing in ("TENSORBOARD_IMAGE", "ISTIO_GATEWAY", "ISTIO_HOST"):
        env = {k: v for k, v in ENV.items() if k != missing}
        outs, st = hostplan.tensorboard(docs, None, env)
        assert (np.asarray(st) == 10).all() and all(o == b"" for o in outs)
    # an invalid RWO_PVC_SCHEDULING is an error only where the reference asks for it: PVC paths
    env = dict(ENV, RWO_PVC_SCHEDULING="yes")
    outs, st = hostplan.tensorboard(docs, None, env)
    assert check(tbs, None, outs, st, env, False) > 5 and (np.asarray(st) == 10).sum() > 5


def test_garbage(hostplan):
    outs, st = hostplan.tensorboard([b"", b"[]", b"{", b'{"metadata":7,"spec":"x"}', b"{}"], None, ENV)
    assert list(st[:3]) == [1, 1, 0] or list(st[:2]) == [1, 1]
    assert st[4] == 0 and json.loads(outs[12])["metadata"] == {"name": ""}


def test_copy_deployment_fields_and_status(wl):
    tb = wl.tensorboard_cr(random.Random(3), 1)
    d = otb.generate_deployment(tb, ENV)
    found = json.loads(json.dumps(d))
    assert otb.copy_deployment_fields(d, found) == (found, False)
    found["spec"]["replicas"] = 3
    found["metadata"]["labels"] = {"stale": "x"}
    to, upd = otb.copy_deployment_fields(d, found)
    assert upd and to["spec"]["replicas"] == 1 and "labels" not in to["metadata"]
    found = json.loads(json.dumps(d))
    found["spec"]["template"]["spec"]["affinity"] = otb.node_affinity("n1")
    to, upd = otb.copy_deployment_fields(d, found)
    assert upd and to["spec"]["template"]["spec"]["affinity"] == {}
    dep = {"status": {"conditions": [{"type": "Available", "lastUpdateTime": "2026-11-07T00:00:00Z"}], "readyReplicas": 1}}
    s = otb.update_status({"status": {"conditions": [{"deploymentState": "Progressing"}], "readyReplicas": 0}}, dep)
    assert s == {"conditions": [{"deploymentState": "Progressing"},
                                {"deploymentState": "Available", "lastProbeTime": "2026-11-07T00:00:00Z"}], "readyReplicas": 1}
    assert otb.update_status({"status": s}, dep) == s and otb.update_status({}, None) is None
    assert otb.update_status({}, {"status": {}}) == {"conditions": None, "readyReplicas": 0}


def deployments_for(wl, tbs, seed):
    """(desired, found) Deployments per Tensorboard: found None (NotFound), equal, or drifted in the
    three things CopyDeploymentSetFields looks at, with what an API server adds."""
    rng = random.Random(seed)
    desired, found = [], []
    for tb in tbs:
        d = otb.generate_deployment(tb, ENV, nodename=rng.choice(["", "node-1"]) if rng.random() < 0.3 else "")
        desired.append(d)
        r = rng.random()
        if r < 0.25:
            found.append(None)
            continue
        f = json.loads(json.dumps(d))
        f["metadata"].update({"resourceVersion": "7", "generation": 3})
        f["status"] = {"readyReplicas": rng.choice([0, 1]), "conditions": rng.choice(
            [[], [{"type": "Available", "status": "True", "lastUpdateTime": "2026-11-07T10:00:00Z"}],
             [{"type": "Progressing", "status": "True"}, {"type": "Available", "status": "True"}]])}
        if rng.random() < 0.2:
            del f["status"]["readyReplicas"]
        if rng.random() < 0.2:
            f["status"] = {}
        f["spec"]["strategy"] = {"type": "RollingUpdate"}
        if r < 0.4:
            f["spec"]["replicas"] = 3
        elif r < 0.55:
            f["metadata"]["labels"] = {"stale": "x"}
        elif r < 0.65:
            f["metadata"]["labels"] = {"empty": ""}          # from.Labels[k] == "" == v: no update
        elif r < 0.8:
            f["spec"]["template"]["spec"]["affinity"] = otb.node_affinity("node-9")
        elif r < 0.87:
            del f["spec"]["template"]["spec"]["affinity"]
        found.append(f)
    return desired, found


def with_status(rng, tb):
    tb = json.loads(json.dumps(tb))
    tb["status"] = rng.choice([None, {}, {"conditions": [], "readyReplicas": 0},
                               {"conditions": [{"deploymentState": "Available", "lastProbeTime": "2026-11-07T09:00:00Z"}],
                                "readyReplicas": 1},
                               {"conditions": [{"deploymentState": "Progressing"}, {"deploymentState": "ReplicaFailure"}]}])
    if tb["status"] is None:
        del tb["status"]
    return tb


def check_update(tbs, desired, found, outs, st, res):
    n_upd = n_same = 0
    for i, tb in enumerate(tbs):
        assert st[i] == 0
        if found[i] is None:
            assert outs[2 * i] == b"" and outs[2 * i + 1] == b"" and res[i] == 0
            continue
        to, upd = otb.copy_deployment_fields(desired[i], found[i])
        assert bool(res[i] & 1) == upd, (desired[i], found[i])
        if upd:
            assert gojson.canon(json.loads(outs[2 * i])) == gojson.canon(to), (outs[2 * i], to)
            n_upd += 1
        else:
            assert outs[2 * i] == b""
            n_same += 1
        assert gojson.canon(json.loads(outs[2 * i + 1])) == gojson.canon(otb.update_status(tb, found[i]))
    return n_upd, n_same


def test_deployment_update_and_status(hostplan, wl):
    rng = random.Random(71)
    tbs = [with_status(rng, wl.tensorboard_cr(rng, i)) for i in range(1500)]
    desired, found = deployments_for(wl, tbs, 72)
    outs, st, res = hostplan.tensorboard_update([wl.dumps(t, rng) for t in tbs], [wl.dumps(d, rng) for d in desired],
                                                [b"" if f is None else wl.dumps(f, rng) for f in found])
    n_upd, n_same = check_update(tbs, desired, found, outs, st, res)
    assert n_upd > 500 and n_same > 150
