"""GPU: what a PVCViewer reconcile writes, through the C ABI (nbrm_pvcviewer_reconcile), against the
oracle and, byte for byte, against the CPU build of the same planner."""
import random

import pytest

from test_hostplan_pvcviewer import ENVS, check, make

pytestmark = pytest.mark.gpu


def test_pvcviewers(ctx, pkg, wl, hostplan):
    viewers, found, nodes = make(wl, 6000, 41)
    rng = random.Random(42)
    docs = [wl.dumps(v, rng) for v in viewers]
    fdocs = [b"" if f is None else wl.dumps(f, rng) for f in found]
    tot = [0, 0, 0]
    try:
        for env in ENVS:
            ctx.set_env(env)
            res = ctx.pvcviewer_reconcile(pkg.Docs(docs), pkg.Docs(fdocs), pkg.Docs(nodes))
            r = check(viewers, found, nodes, res.docs(), res.status, env)
            tot = [a + b for a, b in zip(tot, r)]
            outs, st = hostplan.pvcviewer(docs, fdocs, nodes, env)
            assert outs == res.docs() and list(st) == list(res.status)
            res = ctx.pvcviewer_reconcile(pkg.Docs(docs), None, None)
            check(viewers, None, None, res.docs(), res.status, env)
        assert tot[0] > 8000 and tot[1] > 8000 and tot[2] > 400, tot
        # a small batch (the warp-per-tuple kernel is not built for this unit: same kernel)
        res = ctx.pvcviewer_reconcile(pkg.Docs(docs[:64]), pkg.Docs(fdocs[:192]), pkg.Docs(nodes[:64]))
        check(viewers[:64], found[:192], nodes[:64], res.docs(), res.status, ENVS[-1])
    finally:
        ctx.set_env({})


def test_webhook(ctx, pkg, wl, hostplan):
    import json

    from test_hostplan_pvcviewer import DEFAULT_FILE, check_webhook, webhook_viewers
    viewers = webhook_viewers(wl, 5000, 61)
    rng = random.Random(62)
    docs = [wl.dumps(v, rng) for v in viewers]
    try:
        for f in (None, DEFAULT_FILE):
            env = {} if f is None else {"PVCVIEWER_DEFAULT_PODSPEC_JSON": json.dumps(f)}
            ctx.set_env(env)
            res = ctx.pvcviewer_webhook(pkg.Docs(docs))
            seen = check_webhook(viewers, res.docs(), res.status, res.result, f)
            assert seen[0] > 1500 and seen[1] > 200 and seen[3] > 200
            outs, st, r = hostplan.pvcviewer_webhook(docs, env)
            assert outs == res.docs() and list(r) == list(res.result)
    finally:
        ctx.set_env({})


def test_reconcile_status(ctx, pkg, wl, hostplan):
    from test_hostplan_pvcviewer import check_status, status_cases
    viewers, deps = status_cases(wl, 6000, 93)
    rng = random.Random(94)
    a = [wl.dumps(v, rng) for v in viewers]
    b = [b"" if d is None else wl.dumps(d, rng) for d in deps]
    res = ctx.pvcviewer_status(pkg.Docs(a), pkg.Docs(b))
    check_status(viewers, deps, res.docs(), res.status)
    outs, st = hostplan.pvcviewer_status(a, b)
    assert outs == res.docs()
