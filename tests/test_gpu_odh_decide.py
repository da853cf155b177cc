"""GPU: the ODH write decision through the C ABI, alone and inside the composed pass."""
import importlib
import json
import random

import numpy as np
import pytest

import __graft_entry__ as graft
from oracle import gojson, nbc, odh, passes
from parity import check_odh_decide
from test_hostplan_odh_decide import make_odh_pairs, pack

pytestmark = pytest.mark.gpu
NOW = 1793966400
ENV = {"USE_ISTIO": "true", "CONTROLLER_NAMESPACE": passes.CTRL_NS}


def test_odh_decide_fuzz(ctx, pkg, wl, hostplan):
    desired, found = make_odh_pairs(wl, 1500, 71)
    d, f = pack(wl, desired, found, random.Random(2))
    res = ctx.odh_decide(pkg.Docs(d), pkg.Docs(f))
    c, u, z = check_odh_decide(desired, found, res.docs(), res.status, res.result)
    assert c > 600 and u > 1000 and z > 3000, (c, u, z)
    outs, st, rs = hostplan.odh_decide(d, f)          # same planner source on the CPU: same bytes
    assert outs == res.docs() and list(st) == list(res.status) and list(rs) == list(res.result)


def test_full_pass_with_odh_decision(pkg, wl):
    """The composed pass with the found ODH objects on the device: the "odh" result is the
    reconcilers' write set (oracle: full_pass(..., odh_found=...)), generated objects never
    leave the GPU."""
    rp_mod = importlib.import_module(graft.PKG_NAME + ".reconcile")
    rng = random.Random(5)
    n = 700
    reqs = wl.corpus("c2", n, start=40)
    seeds = rng.randbytes(16 * n)
    now_str = gojson.format_rfc3339(NOW * 10**9)
    found, stss, pods, odh_found, exp = [], [], [], [], []
    for i, r in enumerate(reqs):
        mutated, _ = odh.mutate_notebook(r, "CREATE")
        desired = [nbc.generate_statefulset(mutated, ENV), nbc.generate_service(mutated, ENV),
                   nbc.generate_virtualservice(mutated, ENV)]
        for d in desired:
            nbc.set_controller_reference(mutated, d)
        f = [wl.found_variant(rng, d) for d in desired]
        found += f
        stss.append(f[0])
        pods.append(wl.pod_variant(rng, mutated, NOW))
        cookie = odh.cookie_secret_from_seed(seeds[16 * i:16 * i + 16])
        objs = odh.reconcile_objects(mutated, passes.CTRL_NS, rbac=False, cookie_secret=cookie)
        of = [None if o is None else wl.odh_found_variant(rng, o) for o in objs]
        odh_found += of
        exp.append(passes.full_pass(r, f, f[0], pods[-1], ENV, now_str, cookie, odh_found=of))
    ctx = pkg.Context(0, env=ENV)
    rp = rp_mod.ReconcilePass(ctx, NOW)
    d = lambda xs: pkg.Docs([b"" if x is None else wl.dumps(x) for x in xs])
    rp.set_cluster_state(d(found), d(stss), d(pods), odh_found=d(odh_found))
    for overlap in (False, True):
        out = rp.run(d(reqs), seeds, overlap=overlap)
        odhs, writes = out["odh"].docs(), out["writes"].docs()
        assert out["odh"].n == 7 * n
        wrote = 0
        for i, (mutated, odh_w, w, st) in enumerate(exp):
            for k in range(7):
                got = odhs[7 * i + k]
                assert (got == b"") == (odh_w[k] is None), (i, k, got)
                if odh_w[k] is not None:
                    assert gojson.canon(json.loads(got)) == gojson.canon(odh_w[k]), (i, k)
                    wrote += 1
            for k in range(3):
                got = writes[3 * i + k]
                assert (got == b"") == (w[k] is None), (i, k)
            assert gojson.canon(json.loads(out["status"].doc(i))) == gojson.canon(st), i
        assert wrote > n
    ctx.close()
