"""CPU: the ODH webhook mutation planners (compiled by g++) against the oracle and the
reference's expected OAuth Notebook."""
import copy
import json
import os
import random

from oracle import gojson, odh
from parity import check_mutate

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
IMG = odh.OAUTH_PROXY_IMAGE


def good_notebooks(wl, n, seed):
    """Fuzz Notebooks that have a pod template (the CRD requires containers)."""
    out = []
    for nb in wl.corpus("fuzz", n, start=seed):
        spec = ((nb.get("spec") or {}).get("template") or {}).get("spec")
        if isinstance(spec, dict) and spec.get("containers"):
            out.append(nb)
    return out


def test_create_reference_fixture(hostplan, wl):
    g = json.load(open(os.path.join(G, "odh.json")))["oauth"]
    outs, st = hostplan.mutate([wl.dumps(g["notebook"])], {}, flags=1)
    assert list(st) == [0]
    assert odh.compare_notebooks(json.loads(outs[0]), g["expected_notebook"])


def test_create_fuzz(hostplan, wl):
    nbs = good_notebooks(wl, 2500, 3)
    rng = random.Random(1)
    docs = [wl.dumps(nb, rng) for nb in nbs]
    for flags, ca in ((1, False), (1 | 2, True)):
        outs, st = hostplan.mutate(docs, {"OAUTH_PROXY_IMAGE": "quay.io/x/proxy:1"}, flags)
        ok, denied, blocked = check_mutate(nbs, None, outs, st, [0] * len(nbs), "quay.io/x/proxy:1", ca)
        assert ok > 1800 and denied > 20 and blocked == 0, (ok, denied, blocked)


def update_cases(wl, n, seed):
    """(request, old) pairs for UPDATE: the old object is a created-and-running
    Notebook; the request repeats it, edits metadata, edits the pod template,
    stops or restarts it."""
    rng = random.Random(seed)
    reqs, olds = [], []
    for nb in good_notebooks(wl, n, seed):
        try:
            old, _ = odh.mutate_notebook(nb, "CREATE", image=IMG, ca_bundle=rng.random() < 0.3)
        except odh.Denied:
            continue
        a = old["metadata"].setdefault("annotations", {}) if isinstance(old["metadata"].get("annotations"), dict) \
            else None
        if a is not None:
            a.pop("kubeflow-resource-stopped", None)          # the lock is gone: running
        req = copy.deepcopy(old)
        ra = req["metadata"].get("annotations")
        r = rng.random()
        if r < 0.3:
            pass
        elif r < 0.45 and isinstance(ra, dict):
            ra["opendatahub.io/username"] = "someone"
        elif r < 0.6:
            req["spec"]["template"]["spec"]["containers"][0]["image"] = "user/changed:2"
        elif r < 0.7 and isinstance(ra, dict):
            ra["kubeflow-resource-stopped"] = "2026-11-06T12:00:00Z"
        elif r < 0.8 and isinstance(ra, dict):
            ra["notebooks.opendatahub.io/notebook-restart"] = "true"
        elif r < 0.9 and isinstance(ra, dict):
            ra["notebooks.opendatahub.io/update-pending"] = "stale reason"
        reqs.append(req)
        olds.append(old)
    return reqs, olds


def run_update(hostplan, wl, reqs, olds, env, flags):
    rd, od = [wl.dumps(x) for x in reqs], [wl.dumps(x) for x in olds]
    mut, st1 = hostplan.mutate(rd, env, flags & ~1)
    outs, st, res = hostplan.restart_guard(mut, rd, od, env)
    return outs, st, res


def test_update_fuzz(hostplan, wl):
    reqs, olds = update_cases(wl, 2500, 5)
    # same webhook configuration as at creation: nothing is held back
    outs, st, res = run_update(hostplan, wl, reqs, olds, {}, 0)
    ok, denied, blocked = check_mutate(reqs, olds, outs, st, res, IMG, False)
    assert ok > 1500 and blocked < ok // 2, (ok, denied, blocked)
    # the webhook now injects another proxy image: running OAuth notebooks are guarded
    env = {"OAUTH_PROXY_IMAGE": "registry.example/new-proxy@sha256:abc"}
    outs, st, res = run_update(hostplan, wl, reqs, olds, env, 0)
    ok, denied, blocked = check_mutate(reqs, olds, outs, st, res, env["OAUTH_PROXY_IMAGE"], False)
    assert ok > 1500 and blocked > 100, (ok, denied, blocked)
