"""GPU: status mirroring and the culling pass through the C ABI against the oracle."""
import json
import os
import random

import pytest

from oracle import gojson
from parity import check_cull, check_status
from test_hostplan_status_cull import G, NOW, _d, cull_inputs, pack_cull, status_inputs

pytestmark = pytest.mark.gpu


def test_status_reference_table_and_fuzz(ctx, pkg, wl, hostplan):
    cases = json.load(open(os.path.join(G, "nbc_status.json")))
    res = ctx.notebook_status(pkg.Docs([wl.dumps(c["notebook"]) for c in cases]),
                              pkg.Docs([_d(wl, c["sts"] or None) for c in cases]),
                              pkg.Docs([_d(wl, c["pod"] or None) for c in cases]), NOW)
    for c, o in zip(cases, res.docs()):
        assert gojson.canon(json.loads(o)) == gojson.canon(c["expected"]), c["name"]
    nbs, stss, pods = status_inputs(wl, 5000, 43)
    rng = random.Random(2)
    a, b, d = [wl.dumps(x, rng) for x in nbs], [_d(wl, x, rng) for x in stss], [_d(wl, x, rng) for x in pods]
    res = ctx.notebook_status(pkg.Docs(a), pkg.Docs(b), pkg.Docs(d), NOW)
    assert check_status(nbs, stss, pods, res.docs(), res.status, gojson.format_rfc3339(NOW * 10**9)) == 5000
    outs, st = hostplan.status(a, b, d, now=NOW)
    assert outs == res.docs()


def test_cull_fuzz(ctx, pkg, wl, hostplan):
    cases = cull_inputs(wl, 6000, 19)
    nbs, ks, ts, pe = pack_cull(wl, cases)
    for idle, period in ((5, 1), (1440, 1)):
        res = ctx.culling_pass(pkg.Docs(nbs), pkg.Docs(ks), pkg.Docs(ts), pe, NOW, idle, period)
        seen = check_cull(cases, res.docs(), res.status, res.result, NOW, idle, period)
        assert all(seen[k] > 30 for k in (1, 2, 3, 4)), seen
        outs, st, rs = hostplan.cull(nbs, ks, ts, pe, NOW, idle, period)
        assert outs == res.docs() and list(rs) == list(res.result)
