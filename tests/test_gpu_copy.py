"""GPU: Copy*Fields / create-or-update and the fused reconcile pass through the C ABI."""
import json
import random

import numpy as np
import pytest

from oracle import gojson, nbc
from parity import check_copy_fields
from test_hostplan_copy import ENV, make_pairs, pack_pairs

pytestmark = pytest.mark.gpu


def test_copy_fields_fuzz(ctx, pkg, wl, hostplan):
    desired, found = make_pairs(wl, 3000, 33)
    d, f = pack_pairs(wl, desired, found, random.Random(6))
    res = ctx.copy_fields(pkg.Docs(d), pkg.Docs(f))
    c, u, z = check_copy_fields(desired, found, res.docs(), res.status)
    assert c > 600 and u > 2000 and z > 2000, (c, u, z)
    outs, st = hostplan.copy_fields(d, f)            # same planner source on the CPU: same bytes
    assert outs == res.docs() and list(st) == list(res.status)


def test_fused_reconcile_pass(ctx, pkg, wl):
    """generate on the device -> compare on the device == oracle generate + create_or_update."""
    ctx.set_env(ENV)
    rng = random.Random(8)
    nbs = wl.corpus("fuzz", 2000, start=77)
    desired, found = [], []
    for nb in nbs:
        try:
            objs = [nbc.generate_statefulset(nb, ENV), nbc.generate_service(nb, ENV),
                    nbc.generate_virtualservice(nb, ENV)]
        except nbc.ReferencePanic:
            desired += [None] * 3
            found += [None] * 3
            continue
        for o in objs:
            nbc.set_controller_reference(nb, o)
            desired.append(o)
            found.append(wl.found_variant(rng, o))
    fdocs = [b"" if x is None else wl.dumps(x, rng) for x in found]
    res = ctx.nbc_reconcile(pkg.Docs([wl.dumps(nb) for nb in nbs]), pkg.Docs(fdocs), 7 | 8)
    assert res.n == 3 * len(nbs)
    c, u, z = check_copy_fields(desired, found, res.docs(), res.status, rejected=True)
    assert c > 300 and u > 1000 and z > 1000, (c, u, z)


def test_steady_state_is_silent(ctx, pkg, wl):
    """Found == what the library itself generated (+ server metadata): nothing to write."""
    ctx.set_env(ENV)
    nbs = wl.corpus("c2", 5000)
    docs = pkg.Docs([wl.dumps(nb) for nb in nbs])
    gen = ctx.nbc_generate(docs, 7 | 8)
    res = ctx.nbc_reconcile(docs, pkg.Docs(gen.docs()), 7 | 8)
    assert int(res.off[-1]) == 0 and int(res.status.sum()) == 0
    # and with every object missing, everything is created, byte for byte the generated object
    res = ctx.nbc_reconcile(docs, pkg.Docs([b""] * (3 * len(nbs))), 7 | 8)
    assert np.array_equal(res.data, gen.data) and np.array_equal(res.off, gen.off)


@pytest.mark.parametrize("pad", [0, 3000])
@pytest.mark.parametrize("n", [1, 2, 43, 129, 515])
def test_per_kind_thread_mapping_is_invisible(ctx, pkg, wl, n, pad):
    """The fused pass knows its desired objects repeat with period 3 and gives the threads
    that run together one kind each (plan_kernel: per warp when staged, per block when the
    tuples are too large to stage — pad makes them so).  Results must equal the same compare
    on uploaded documents, where no period is known, at sizes around the block edges."""
    ctx.set_env(ENV)
    rng = random.Random(1000 + n)
    nbs = wl.corpus("fuzz", n, start=500 + n)
    for nb in nbs:
        if pad:
            nb["metadata"].setdefault("annotations", {})["pad"] = "x" * pad
    docs = pkg.Docs([wl.dumps(nb) for nb in nbs])
    gen = ctx.nbc_generate(docs, 7 | 8)
    gdocs = gen.docs()
    found = []
    for g in gdocs:
        if not g or rng.random() < 0.2:
            found.append(b"")
        else:
            found.append(wl.dumps(wl.found_variant(rng, json.loads(g)), rng))
    plain = ctx.copy_fields(pkg.Docs(gdocs), pkg.Docs(found))
    p_docs, p_status = plain.docs(), list(plain.status)
    fused = ctx.nbc_reconcile(docs, pkg.Docs(found), 7 | 8)
    assert fused.n == 3 * n
    assert fused.docs() == p_docs
    # a Notebook the generators rejected carries their status on its three pairs in the fused
    # pass; the plain compare only ever saw three empty desired documents for it
    gen_status = [int(gen.status[i // 3]) for i in range(3 * n)]
    assert list(fused.status) == [g or p for g, p in zip(gen_status, p_status)]
