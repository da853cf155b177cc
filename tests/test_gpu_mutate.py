"""GPU: the ODH webhook mutation through the C ABI against the oracle."""
import json
import os
import random

import pytest

from oracle import odh
from parity import check_mutate
from test_hostplan_mutate import G, IMG, good_notebooks, update_cases

pytestmark = pytest.mark.gpu


def test_create(ctx, pkg, wl, hostplan):
    g = json.load(open(os.path.join(G, "odh.json")))["oauth"]
    ctx.set_env({})
    res = ctx.odh_mutate(pkg.Docs([wl.dumps(g["notebook"])]))
    assert odh.compare_notebooks(json.loads(res.doc(0)), g["expected_notebook"])
    nbs = good_notebooks(wl, 4000, 9)
    rng = random.Random(1)
    docs = [wl.dumps(nb, rng) for nb in nbs]
    env = {"OAUTH_PROXY_IMAGE": "quay.io/x/proxy:1"}
    ctx.set_env(env)
    for flags, ca in ((0, False), (pkg.MUT_CA_BUNDLE, True)):
        res = ctx.odh_mutate(pkg.Docs(docs), None, flags)
        ok, denied, blocked = check_mutate(nbs, None, res.docs(), res.status, res.result, env["OAUTH_PROXY_IMAGE"], ca)
        assert ok > 3000 and denied > 30 and blocked == 0
        outs, st = hostplan.mutate(docs, env, flags | 1)
        assert outs == res.docs() and list(st) == list(res.status)


def test_update(ctx, pkg, wl):
    reqs, olds = update_cases(wl, 4000, 15)
    rd, od = pkg.Docs([wl.dumps(x) for x in reqs]), pkg.Docs([wl.dumps(x) for x in olds])
    ctx.set_env({})
    res = ctx.odh_mutate(rd, od, 0)
    ok, denied, blocked = check_mutate(reqs, olds, res.docs(), res.status, res.result, IMG, False)
    assert ok > 2500
    env = {"OAUTH_PROXY_IMAGE": "registry.example/new-proxy@sha256:abc"}
    ctx.set_env(env)
    res = ctx.odh_mutate(rd, od, 0)
    ok, denied, blocked = check_mutate(reqs, olds, res.docs(), res.status, res.result, env["OAUTH_PROXY_IMAGE"], False)
    assert ok > 2500 and blocked > 200
    ctx.set_env({})
