"""GPU: the tensorboard-controller's desired objects through the C ABI (nbrm_tensorboard_generate)
against the oracle and, byte for byte, against the CPU build of the same planner."""
import random

import numpy as np
import pytest

from test_hostplan_tensorboard import ENV, ENVS, check

pytestmark = pytest.mark.gpu


def test_tensorboards(ctx, pkg, wl, hostplan):
    rng = random.Random(21)
    tbs = [wl.tensorboard_cr(rng, i) for i in range(3000)]
    docs = [wl.dumps(t, rng) for t in tbs]
    nodes = [wl.tensorboard_node(rng) for _ in tbs]
    seen = 0
    try:
        for env in ENVS:
            ctx.set_env(env)
            for flags in (0, 8):
                res = ctx.tensorboard_generate(pkg.Docs(docs), pkg.Docs(nodes), flags)
                seen += check(tbs, nodes, res.docs(), res.status, env, bool(flags))
                outs, st = hostplan.tensorboard(docs, nodes, env, flags=flags)
                assert outs == res.docs() and list(st) == list(res.status)
            res = ctx.tensorboard_generate(pkg.Docs(docs), None, 8)
            check(tbs, None, res.docs(), res.status, env, True)
        assert seen > 20000
        # small batch: the warp-per-tuple kernel
        res = ctx.tensorboard_generate(pkg.Docs(docs[:100]), pkg.Docs(nodes[:100]), 8)
        assert check(tbs[:100], nodes[:100], res.docs(), res.status, ENVS[-1], True) > 50
        # a required variable is not set
        ctx.set_env({k: v for k, v in ENV.items() if k != "ISTIO_HOST"})
        res = ctx.tensorboard_generate(pkg.Docs(docs[:50]), None, 0)
        assert (np.asarray(res.status) == 10).all() and all(o == b"" for o in res.docs())
    finally:
        ctx.set_env({})


def test_deployment_update_and_status(ctx, pkg, wl, hostplan):
    from test_hostplan_tensorboard import check_update, deployments_for, with_status
    rng = random.Random(91)
    tbs = [with_status(rng, wl.tensorboard_cr(rng, i)) for i in range(6000)]
    desired, found = deployments_for(wl, tbs, 92)
    a = [wl.dumps(t, rng) for t in tbs]
    b = [wl.dumps(d, rng) for d in desired]
    c = [b"" if f is None else wl.dumps(f, rng) for f in found]
    res = ctx.tensorboard_update(pkg.Docs(a), pkg.Docs(b), pkg.Docs(c))
    n_upd, n_same = check_update(tbs, desired, found, res.docs(), res.status, res.result)
    assert n_upd > 2000 and n_same > 600
    outs, st, r = hostplan.tensorboard_update(a, b, c)
    assert outs == res.docs() and list(r) == list(res.result)
