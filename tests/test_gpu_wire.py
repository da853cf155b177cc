"""GPU: JSON patches and AdmissionReview responses through the C ABI (nbrm_json_diff,
nbrm_notebook_webhook_review, nbrm_poddefault_review) against the oracle (oracle/wire.py) and,
byte for byte, against the same planner sources compiled by g++."""
import copy
import json
import random

import pytest

from oracle import wire
from test_hostplan_wire import _mutations, _no_nulls

pytestmark = pytest.mark.gpu


def test_json_diff(ctx, pkg, wl, hostplan):
    from oracle import gojson, poddefault
    rng = random.Random(9)
    origs = wl.corpus("fuzz", 1500, start=21) + wl.corpus("c2", 30)
    muts = [_mutations(rng, nb) if rng.random() < 0.9 else copy.deepcopy(nb) for nb in origs]
    a = [wl.dumps(x, rng) for x in origs]
    b = [wl.dumps(x, rng) for x in muts]
    res = ctx.json_diff(pkg.Docs(a), pkg.Docs(b))
    outs, st = hostplan.json_diff(a, b)
    assert res.docs() == outs and list(res.status) == list(st) and not any(st)
    n = 0
    for i, (o, m) in enumerate(zip(origs, muts)):
        if not outs[i]:
            assert gojson.canon(_no_nulls(o)) == gojson.canon(_no_nulls(m))
            continue
        got = poddefault.apply_json_patch(json.loads(a[i]), json.loads(outs[i]))
        assert gojson.canon(_no_nulls(got)) == gojson.canon(_no_nulls(m)), i
        n += 1
    assert n > 1200


def test_notebook_webhook_reviews(ctx, pkg, wl):
    ctx.set_env({})
    nbs = [nb for nb in wl.corpus("fuzz", 1200, start=5)
           if isinstance(((nb.get("spec") or {}).get("template") or {}).get("spec"), dict)
           and nb["spec"]["template"]["spec"].get("containers")] + wl.corpus("c2", 50)
    reviews = [wire.review_request(nb, "uid-%d" % i, resource={"group": "kubeflow.org", "version": "v1", "resource": "notebooks"})
               for i, nb in enumerate(nbs)]
    rng = random.Random(4)
    res = ctx.notebook_webhook_review(pkg.Docs([wl.dumps(r, rng) for r in reviews]))
    seen = {"denied": 0, "patched": 0}
    for i, (r, raw) in enumerate(zip(reviews, res.docs())):
        assert res.status[i] == 0
        body = json.loads(raw)
        assert body["kind"] == "AdmissionReview" and body["apiVersion"] == "admission.k8s.io/v1"
        uid, allowed, patch, status, has = wire.decode_response(raw)
        exp = wire.notebook_webhook_response(r)
        assert has and uid == "uid-%d" % i
        if exp[0] == "denied":
            assert allowed is False and patch is None and status["code"] == 403 and status["message"] == exp[1]
            seen["denied"] += 1
        else:
            assert allowed is True and patch                      # CREATE always adds the lock
            wire.check_patch_effect(_no_nulls(r["request"]["object"]), patch, _no_nulls(exp[1]))
            seen["patched"] += 1
    assert seen["denied"] > 5 and seen["patched"] > 800, seen


def test_poddefault_reviews(ctx, pkg, wl):
    pods, pds, ranges = wl.admission_workload(1000, 10)
    rng = random.Random(2)
    reviews = []
    for i, pod in enumerate(pods):
        p = copy.deepcopy(pod)
        ns = p["metadata"]["namespace"]
        if rng.random() < 0.3:
            del p["metadata"]["namespace"]
        other = {"group": "apps", "version": "v1", "resource": "deployments"} if rng.random() < 0.1 else None
        reviews.append(wire.review_request(p, "uid-%d" % i, namespace=ns, resource=other))
    res = ctx.poddefault_review(pkg.Docs([wl.dumps(r, rng) for r in reviews]), pkg.Docs([wl.dumps(p) for p in pds]), ranges)
    seen = {"empty": 0, "patched": 0, "plain": 0, "error": 0}
    for i, (r, raw) in enumerate(zip(reviews, res.docs())):
        lo, hi = ranges[i]
        exp = wire.poddefault_response(r, pds[lo:hi])
        uid, allowed, patch, status, has = wire.decode_response(raw)
        if exp[0] == "empty":
            assert not has and json.loads(raw) == {}
            seen["empty"] += 1
            continue
        assert has and uid == "uid-%d" % i
        if exp[0] == "error":
            assert allowed is False and patch is None and status["message"]
            seen["error"] += 1
        elif exp[1] is None:
            assert allowed is True and patch is None and status is None
            seen["plain"] += 1
        else:
            pod = copy.deepcopy(r["request"]["object"])
            pod["metadata"].setdefault("namespace", r["request"]["namespace"])
            assert allowed is True
            wire.check_patch_effect(pod, patch, exp[1])
            seen["patched"] += 1
    assert seen["empty"] > 50 and seen["patched"] > 100 and seen["plain"] > 100 and seen["error"] > 0, seen


def test_notebook_version_conversion(ctx, pkg, wl, hostplan):
    """nbrm_notebook_convert against the oracle's ConvertTo / ConvertFrom and the CPU build."""
    from oracle import gojson, nbc
    from test_hostplan_convert import samples, with_status
    rng = random.Random(4)
    nbs = [with_status(rng, nb) for nb in list(samples().values()) + wl.corpus("fuzz", 2000, start=9)]
    docs = [wl.dumps(nb, rng) for nb in nbs]
    for to_version, flags in ((0, 0), (1, 1), (2, 0)):
        res = ctx.notebook_convert(pkg.Docs(docs), to_version, flags)
        outs, st = hostplan.convert(docs, to_version, flags)
        assert res.docs() == outs and not any(res.status)
        for nb, o in zip(nbs, outs):
            want = nbc.convert_notebook(nb, nbc.API_VERSIONS[to_version], keep_metadata=bool(flags))
            assert gojson.canon(json.loads(o)) == gojson.canon(want)
