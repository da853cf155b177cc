"""GPU: NBRM_OPT_GROUP_BY_LENGTH changes which thread plans which tuple and nothing else: every
generator of the reconcile pass gives byte-identical results, statuses and result words with the
option on and off, on a batch that mixes shapes and on a uniform one."""
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def run_all(ctx, pkg, wl, nbs, docs):
    out = []
    m = ctx.odh_mutate(pkg.Docs(docs))
    out.append((m.docs(), list(m.status)))
    mutated = [d if d else b"{}" for d in m.docs()]
    r = ctx.nbc_generate(pkg.Docs(mutated), flags=pkg.GEN_ALL | pkg.GEN_OWNER_REF | pkg.GEN_OWNER_V1)
    out.append((r.docs(), list(r.status)))
    r = ctx.odh_generate(pkg.Docs(mutated), seeds=bytes(16 * len(mutated)), flags=pkg.GEN_OWNER_REF | pkg.GEN_OWNER_V1)
    out.append((r.docs(), list(r.status)))
    return out


def test_identical_results(pkg, wl):
    rng = random.Random(5)
    nbs = wl.corpus("fuzz", 3000, start=11) + wl.corpus("c2", 2000) + wl.corpus("c1", 1500) + wl.corpus("loadtest", 500)
    rng.shuffle(nbs)
    docs = [wl.dumps(nb, rng) for nb in nbs]
    assert len(docs) > 4096      # above the warp-kernel threshold: the thread kernel, where the option acts
    ctx = pkg.Context(0, env={"USE_ISTIO": "true"})
    try:
        base = run_all(ctx, pkg, wl, nbs, docs)
        ctx.set_option(pkg.capi.OPT_GROUP_BY_LENGTH, 1)
        grouped = run_all(ctx, pkg, wl, nbs, docs)
        ctx.set_option(pkg.capi.OPT_GROUP_BY_LENGTH, 0)
        again = run_all(ctx, pkg, wl, nbs, docs)
        for a, b, c in zip(base, grouped, again):
            assert a[0] == b[0] == c[0] and a[1] == b[1] == c[1]
        assert sum(len(x) > 0 for x in base[1][0]) > 10000
    finally:
        ctx.close()
