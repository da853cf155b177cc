"""GPU: ODH reconcile generators through the C ABI against the oracle."""
import random

import pytest

from parity import check_odh_generate
from test_hostplan_odh import CTRL_NS, _seeds

pytestmark = pytest.mark.gpu


def test_fuzz_corpus(ctx, pkg, wl):
    ctx.set_env({"CONTROLLER_NAMESPACE": CTRL_NS})
    nbs = wl.corpus("fuzz", 4000, start=11)
    rng = random.Random(5)
    docs = pkg.Docs([wl.dumps(nb, rng) for nb in nbs])
    seeds = _seeds(len(nbs), 2)
    for flags in (8 | 16, 8 | 16 | 32, 0):
        res = ctx.odh_generate(docs, seeds, flags)
        assert check_odh_generate(nbs, res.docs(), res.status, CTRL_NS, seeds, flags) > 12000


def test_matches_cpu_build_of_planner_bytes(ctx, pkg, wl, hostplan):
    env = {"CONTROLLER_NAMESPACE": CTRL_NS}
    ctx.set_env(env)
    nbs = wl.corpus("fuzz", 1500, start=5) + wl.corpus("c2", 500)
    docs = [wl.dumps(nb) for nb in nbs]
    seeds = _seeds(len(nbs), 9)
    res = ctx.odh_generate(pkg.Docs(docs), seeds, 8 | 16 | 32)
    outs, st = hostplan.odh_generate(docs, seeds, env, 8 | 16 | 32)
    assert list(st) == list(res.status) and outs == res.docs()
    # no seeds: the all-zero seed (tests only)
    res = ctx.odh_generate(pkg.Docs(docs[:10]), None, 8 | 16)
    outs, st = hostplan.odh_generate(docs[:10], None, env, 8 | 16)
    assert outs == res.docs()
    res = ctx.odh_generate(pkg.Docs([]), None, 8)
    assert res.n == 0
