"""CPU: RFC 6902 patches (GenJsonDiff) and the AdmissionReview envelope (GenReviewOpen /
GenReviewClose) of kubeflow-rm_b200/csrc/nbrm_patch.cuh, compiled by g++, against the oracle
(oracle/wire.py).  Patches are compared by effect."""
import base64
import copy
import json
import random

from oracle import gojson, odh, poddefault, wire


def _mutations(rng, tree):
    """A mutated copy of `tree`: members added, removed, replaced, arrays grown or rewritten."""
    t = copy.deepcopy(tree)

    def walk(x, depth):
        if isinstance(x, dict):
            for k in list(x):
                r = rng.random()
                if r < 0.08:
                    del x[k]
                elif r < 0.16:
                    x[k] = rng.choice([1, "s", {"n": [1, 2]}, [], None, True])
                else:
                    walk(x[k], depth + 1)
            if rng.random() < 0.3:
                x[rng.choice(["new", "a/b", "t~x", "kubeflow-resource-stopped", "x.y/z~w"])] = rng.choice(
                    ["v", {"k": "v"}, [1], 0])
        elif isinstance(x, list):
            r = rng.random()
            if r < 0.25:
                x.append(rng.choice([{"name": "added"}, "s", 3]))
                if rng.random() < 0.5:
                    x.append({"name": "added-2", "deep": {"er": [1]}})
            elif r < 0.35 and x:
                x.pop(rng.randrange(len(x)))
            elif r < 0.45 and x:
                x.insert(0, "front")
            else:
                for v in x:
                    walk(v, depth + 1)
    walk(t, 0)
    return t


def _no_nulls(x):
    """null members count as absent on both sides (Go nil), as everywhere in the library."""
    if isinstance(x, dict):
        return {k: _no_nulls(v) for k, v in x.items() if v is not None}
    if isinstance(x, list):
        return [_no_nulls(v) for v in x]
    return x


def test_json_diff_by_effect(hostplan, wl):
    rng = random.Random(5)
    origs, muts = [], []
    for nb in wl.corpus("fuzz", 600, start=11) + wl.corpus("c2", 20):
        origs.append(nb)
        muts.append(_mutations(rng, nb) if rng.random() < 0.9 else copy.deepcopy(nb))
    a = [wl.dumps(x, rng) for x in origs]
    b = [wl.dumps(x, rng) for x in muts]
    outs, st = hostplan.json_diff(a, b)
    n_ops = 0
    for i, (o, m) in enumerate(zip(origs, muts)):
        assert st[i] == 0, (i, st[i])
        if gojson.canon(_no_nulls(o)) == gojson.canon(_no_nulls(m)):
            assert outs[i] == b"", (i, outs[i])
            continue
        patch = json.loads(outs[i])
        assert patch and all(op["op"] in ("add", "remove", "replace") for op in patch)
        n_ops += len(patch)
        got = poddefault.apply_json_patch(json.loads(a[i]), patch)
        assert gojson.canon(_no_nulls(got)) == gojson.canon(_no_nulls(m)), (i, outs[i])
    assert n_ops > 1500


def test_notebook_webhook_patch_is_small_and_right(hostplan, wl):
    """mutate -> diff: the patch that turns the submitted Notebook into the mutated one adds the
    lock annotation, the sidecar, the two volumes and the service account — not the whole object."""
    nbs = [nb for nb in wl.corpus("c2", 50)]
    docs = [wl.dumps(nb) for nb in nbs]
    mutated, st = hostplan.mutate(docs, {}, flags=1)
    assert not any(st)
    outs, st = hostplan.json_diff(docs, mutated)
    for nb, m, p in zip(nbs, mutated, outs):
        patch = json.loads(p)
        exp, _ = odh.mutate_notebook(nb, "CREATE")
        wire.check_patch_effect(nb, patch, exp)
        paths = sorted(op["path"] for op in patch)
        assert paths == ["/metadata/annotations/kubeflow-resource-stopped", "/spec/template/spec/containers/-",
                         "/spec/template/spec/serviceAccountName", "/spec/template/spec/volumes/-",
                         "/spec/template/spec/volumes/-"], paths
        assert len(p) < len(m)              # the sidecar is most of either; the rest of the Notebook stays home


def test_admission_review_round_trip(hostplan, wl):
    """open -> PodDefault merge -> close: responses as serve writes them (main.go:708-749)."""
    pods, pds, ranges = wl.admission_workload(300, 10)
    rng = random.Random(2)
    reviews = []
    for i, pod in enumerate(pods):
        p = copy.deepcopy(pod)
        ns = p["metadata"]["namespace"]
        if rng.random() < 0.3:
            del p["metadata"]["namespace"]          # filled from request.namespace (:617-620)
        res = None
        if rng.random() < 0.1:
            res = {"group": "apps", "version": "v1", "resource": "deployments"}
        reviews.append(wire.review_request(p, "uid-%d" % i, namespace=ns, resource=res))
    rdocs = [wl.dumps(r, rng) for r in reviews]
    objs, st, bits = hostplan.review_open(rdocs)
    assert not any(st)
    for r, o, b in zip(reviews, objs, bits):
        want = copy.deepcopy(r["request"]["object"])
        want["metadata"].setdefault("namespace", r["request"]["namespace"])
        assert gojson.canon(json.loads(o)) == gojson.canon(want)
        assert bool(b & 1) == (r["request"]["resource"]["resource"] == "pods") and b & 2
    patches, pst, res = hostplan.poddefault(objs, [wl.dumps(p) for p in pds], ranges)
    outs, cst = hostplan.review_close(rdocs, patches, pst, flags=4)
    seen = {"empty": 0, "patched": 0, "plain": 0, "error": 0}
    for i, (r, raw) in enumerate(zip(reviews, outs)):
        assert cst[i] == 0
        lo, hi = ranges[i]
        exp = wire.poddefault_response(r, pds[lo:hi])
        uid, allowed, patch, status, has = wire.decode_response(raw)
        if exp[0] == "empty":
            assert not has and json.loads(raw) == {}
            seen["empty"] += 1
            continue
        assert has and uid == r["request"]["uid"]
        if exp[0] == "error":
            assert allowed is False and patch is None and status["message"]
            seen["error"] += 1
        elif exp[1] is None:
            assert allowed is True and patch is None and status is None
            seen["plain"] += 1
        else:
            assert allowed is True
            wire.check_patch_effect(json.loads(objs[i]), patch, exp[1])
            seen["patched"] += 1
    assert seen["empty"] > 10 and seen["patched"] > 30 and seen["plain"] > 30 and seen["error"] > 0, seen


def test_notebook_review_controller_runtime_style(hostplan, wl):
    nbs = wl.corpus("c2", 30)
    bad = wl.notebook_c2(99)
    bad["metadata"]["annotations"]["opendatahub.io/service-mesh"] = "true"   # with inject-oauth: denied
    nbs.append(bad)
    reviews = [wire.review_request(nb, "u%d" % i, resource={"group": "kubeflow.org", "version": "v1", "resource": "notebooks"})
               for i, nb in enumerate(nbs)]
    rdocs = [wl.dumps(r) for r in reviews]
    objs, st, bits = hostplan.review_open(rdocs)
    mutated, mst = hostplan.mutate(objs, {}, flags=1)
    assert list(mst[:-1]) == [0] * 30 and mst[-1] == 6
    patches, dst = hostplan.json_diff(objs, mutated)
    outs, cst = hostplan.review_close(rdocs, patches, mst, flags=1)
    for i, (r, raw) in enumerate(zip(reviews, outs)):
        body = json.loads(raw)
        assert body["kind"] == "AdmissionReview" and body["apiVersion"] == "admission.k8s.io/v1"
        uid, allowed, patch, status, has = wire.decode_response(raw)
        exp = wire.notebook_webhook_response(r)
        assert uid == "u%d" % i
        if exp[0] == "denied":
            assert allowed is False and status["code"] == 403 and status["message"] == exp[1]
        else:
            assert allowed is True
            wire.check_patch_effect(r["request"]["object"], patch, exp[1])
