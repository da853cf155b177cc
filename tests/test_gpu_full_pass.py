"""GPU: the composed reconcile pass (webhook -> ODH objects -> notebook-controller objects ->
create-or-update -> status) against the oracle's full_pass, objects device-resident between stages."""
import importlib
import json
import random

import numpy as np
import pytest

import __graft_entry__ as graft
from oracle import gojson, odh, passes

pytestmark = pytest.mark.gpu
NOW = 1793966400
ENV = {"USE_ISTIO": "true", "CONTROLLER_NAMESPACE": passes.CTRL_NS}


def test_full_pass_matches_oracle(pkg, wl):
    rp_mod = importlib.import_module(graft.PKG_NAME + ".reconcile")
    rng = random.Random(12)
    n = 1500
    reqs = wl.corpus("c2", n // 2) + [nb for nb in wl.corpus("fuzz", n, start=500)
                                       if ((nb.get("spec") or {}).get("template") or {}).get("spec", {}).get("containers")
                                       and not (odh.oauth_injection_is_enabled(nb["metadata"]) and
                                                odh.service_mesh_is_enabled(nb["metadata"]))][: n // 2]
    seeds = rng.randbytes(16 * len(reqs))
    now_str = gojson.format_rfc3339(NOW * 10**9)
    found, stss, pods, exp = [], [], [], []
    for i, r in enumerate(reqs):
        mutated, _ = odh.mutate_notebook(r, "CREATE")
        from oracle import nbc
        try:
            desired = [nbc.generate_statefulset(mutated, ENV), nbc.generate_service(mutated, ENV),
                       nbc.generate_virtualservice(mutated, ENV)]
        except nbc.ReferencePanic:
            desired = None
        f = [None, None, None]
        if desired is not None:
            for d in desired:
                nbc.set_controller_reference(mutated, d)
            f = [wl.found_variant(rng, d) for d in desired]
        found += f
        stss.append(f[0])
        pods.append(wl.pod_variant(rng, mutated, NOW))
        cookie = odh.cookie_secret_from_seed(seeds[16 * i:16 * i + 16])
        exp.append(None if desired is None else
                   passes.full_pass(r, f, f[0], pods[-1], ENV, now_str, cookie))
    ctx = pkg.Context(0, env=ENV)
    rp = rp_mod.ReconcilePass(ctx, NOW)
    d = lambda xs: pkg.Docs([b"" if x is None else wl.dumps(x) for x in xs])
    rp.set_cluster_state(d(found), d(stss), d(pods))
    out = rp.run(d(reqs), seeds)
    nbs, odhs, writes, status = (out[k].docs() for k in ("notebooks", "odh", "writes", "status"))
    checked = 0
    for i, e in enumerate(exp):
        if e is None:
            continue
        mutated, odh_objs, w, st = e
        assert gojson.canon(json.loads(nbs[i])) == gojson.canon(mutated), i
        for k in range(7):
            got = odhs[7 * i + k]
            assert (got == b"") == (odh_objs[k] is None), (i, k)
            if odh_objs[k] is not None:
                assert gojson.canon(json.loads(got)) == gojson.canon(odh_objs[k]), (i, k)
        for k in range(3):
            got = writes[3 * i + k]
            assert (got == b"") == (w[k] is None), (i, k)
            if w[k] is not None:
                assert gojson.canon(json.loads(got)) == gojson.canon(w[k]), (i, k)
        assert gojson.canon(json.loads(status[i])) == gojson.canon(st), i
        checked += 1
    assert checked > 1000
    # the overlapped downloads (one lane per result, copies on a second stream while the
    # next stage runs) deliver the same bytes; three passes so the buffers rotate fully
    for _ in range(3):
        again = rp.run(d(reqs), seeds, overlap=True)
        for k in ("notebooks", "odh", "writes", "status"):
            assert np.array_equal(again[k].off, out[k].off), k
            assert np.array_equal(again[k].data, out[k].data), k
            assert np.array_equal(again[k].status, out[k].status), k
            assert np.array_equal(again[k].result, out[k].result), k
    # pipelined batches (requests of batch i+1 uploaded on a stream of their own while batch i
    # runs): three different batches, same bytes as one run() per batch
    third = len(reqs) // 3
    parts = [(pkg.Docs([wl.dumps(x) for x in reqs[k * third:(k + 1) * third]], pinned=True),
              seeds[16 * k * third:16 * (k + 1) * third]) for k in range(3)]
    rp2 = rp_mod.ReconcilePass(ctx, NOW)
    for k in range(3):
        lo, hi = k * third, (k + 1) * third
        rp2.set_cluster_state(d(found[3 * lo:3 * hi]), d(stss[lo:hi]), d(pods[lo:hi]))
        want = rp2.run(parts[k][0], parts[k][1])
        got = list(rp2.run_pipelined([parts[k], parts[k]], copy=True))
        for g in got:
            for name in ("notebooks", "odh", "writes", "status"):
                assert np.array_equal(g[name].off, want[name].off), (k, name)
                assert np.array_equal(g[name].data, want[name].data), (k, name)
                assert np.array_equal(g[name].status, want[name].status), (k, name)
    rp.set_cluster_state(d(found), d(stss), d(pods))
    rp.run(d(reqs), seeds)
    # a lane holds one download at a time
    ctx.run(pkg.G_STATUS, rp_mod.SLOT_NOTEBOOKS, rp_mod.SLOT_STS, rp_mod.SLOT_PODS, now=NOW)
    ctx.download_begin(0)
    ctx.run(pkg.G_STATUS, rp_mod.SLOT_NOTEBOOKS, rp_mod.SLOT_STS, rp_mod.SLOT_PODS, now=NOW)
    with pytest.raises(pkg.NbrmError):
        ctx.download_begin(0)
    st = ctx.download_wait(0)
    assert np.array_equal(st.data, out["status"].data)
    with pytest.raises(pkg.NbrmError):
        ctx.download_wait(0)
    ctx.close()
