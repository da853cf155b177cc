"""CPU: status mirroring and the culling pass (planners compiled by g++) against the oracle
and the reference's own tables."""
import json
import os
import random

from oracle import gojson
from parity import check_cull, check_status

NOW = 1793966400            # 2026-11-06T12:00:00Z
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _d(wl, x, rng=None):
    return b"" if x is None else wl.dumps(x, rng)


def status_inputs(wl, n, seed):
    rng = random.Random(seed)
    nbs, stss, pods = [], [], []
    for nb in wl.corpus("fuzz", n, start=seed):
        nb = wl.status_notebook(rng, nb)
        nbs.append(nb)
        stss.append(wl.sts_variant(rng, nb))
        pods.append(wl.pod_variant(rng, nb, NOW))
    return nbs, stss, pods


def test_status_reference_table(hostplan, wl):
    cases = json.load(open(os.path.join(G, "nbc_status.json")))
    nbs = [c["notebook"] for c in cases]
    outs, st = hostplan.status([wl.dumps(c["notebook"]) for c in cases],
                               [_d(wl, c["sts"] or None) for c in cases],
                               [_d(wl, c["pod"] or None) for c in cases], now=NOW)
    for c, o in zip(cases, outs):
        assert gojson.canon(json.loads(o)) == gojson.canon(c["expected"]), c["name"]


def test_status_fuzz(hostplan, wl):
    nbs, stss, pods = status_inputs(wl, 2500, 41)
    rng = random.Random(2)
    outs, st = hostplan.status([wl.dumps(x, rng) for x in nbs], [_d(wl, x, rng) for x in stss],
                               [_d(wl, x, rng) for x in pods], now=NOW)
    assert check_status(nbs, stss, pods, outs, st, gojson.format_rfc3339(NOW * 10**9)) == 2500


def cull_inputs(wl, n, seed):
    rng = random.Random(seed)
    return [wl.culling_case(rng, i, NOW) for i in range(n)]


def pack_cull(wl, cases):
    nbs = [wl.dumps(c[0]) for c in cases]
    ks = [b"" if c[2] is None else json.dumps(c[2]).encode() for c in cases]
    ts = [b"" if c[3] is None else json.dumps(c[3], indent=1).encode() for c in cases]
    pe = bytes(1 if c[1] else 0 for c in cases)
    return nbs, ks, ts, pe


def test_cull_fuzz(hostplan, wl):
    cases = cull_inputs(wl, 4000, 17)
    nbs, ks, ts, pe = pack_cull(wl, cases)
    for idle, period in ((5, 1), (1440, 1), (0, 0)):
        outs, st, res = hostplan.cull(nbs, ks, ts, pe, NOW, idle, period)
        seen = check_cull(cases, outs, st, res, NOW, idle, period)
        assert all(seen[k] > 20 for k in (1, 2, 3, 4)), seen
        if idle < 1440:
            assert seen[5] > 100, seen


def test_cull_reference_table(hostplan, wl):
    """notebookIsIdle cases of culling_controller_test.go:143-263, as full passes."""
    g = json.load(open(os.path.join(G, "culling.json")))
    for c in g["notebook_is_idle"]["cases"]:
        idle = int(c["env"].get("CULL_IDLE_TIME", 5 if "deadline" in c["name"] or "current" in c["name"] else 1440))
        meta = json.loads(json.dumps(c["meta"]))
        ann = meta.get("annotations")
        if ann is None or "notebooks.kubeflow.org/last-activity" not in ann or "kubeflow-resource-stopped" in ann:
            continue
        for k, v in ann.items():
            if v.startswith("now"):
                ann[k] = gojson.format_rfc3339((NOW + int(v[3:] or 0)) * 10**9)
        ann["notebooks.kubeflow.org/last_activity_check_timestamp"] = gojson.format_rfc3339((NOW - 3600) * 10**9)
        nb = {"metadata": dict(meta, name="n", namespace="ns")}
        outs, st, res = hostplan.cull([wl.dumps(nb)], [b""], [b""], b"\x01", NOW, idle, 1)
        culled = (int(res[0]) & 0xff) == 5
        assert culled == c["result"], c["name"]
