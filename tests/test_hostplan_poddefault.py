"""CPU: the PodDefault admission planner (compiled by g++) against the oracle."""
import json
import os

from oracle import poddefault
from parity import check_poddefault

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def pack(wl, pods, pds):
    return [wl.dumps(p) for p in pods], [wl.dumps(p) for p in pds]


def test_reference_cases(hostplan, wl):
    """admission-webhook/main_test.go:80-252: the four apply cases give no conflict and
    the marker annotation; :12-28 the conflicting annotation is rejected."""
    g = json.load(open(os.path.join(G, "adm.json")))
    for c in g["apply_no_conflict"]["cases"]:
        pod = json.loads(json.dumps(c["pod"]))
        pod.setdefault("metadata", {})
        pds = [dict(p, metadata={"name": "", "namespace": ""}) for p in c["poddefaults"]]
        for p in pds:
            p["spec"]["selector"] = {}
        outs, st, res = hostplan.poddefault([wl.dumps(pod)], [wl.dumps(p) for p in pds])
        seen = check_poddefault([pod], pds, None, outs, st, res)
        assert seen["patched"] == 1, c["name"]
    bad = g["merge_map_bad"]["cases"][0]
    pod = {"metadata": {"annotations": bad["existing"]}, "spec": {}}
    pd = {"metadata": {"name": "x"}, "spec": {"selector": {}, "annotations": bad["defaults"]}}
    outs, st, res = hostplan.poddefault([wl.dumps(pod)], [wl.dumps(pd)])
    assert list(st) == [7] and outs == [b""]


def test_c3_workload(hostplan, wl):
    for per_ns in (1, 10, 100):
        pods, pds, ranges = wl.admission_workload(600, per_ns)
        pd_docs = pack(wl, pods, pds)
        outs, st, res = hostplan.poddefault(pd_docs[0], pd_docs[1], ranges)
        seen = check_poddefault(pods, pds, ranges, outs, st, res)
        assert seen["patched"] > 30 and seen["none"] > 30, (per_ns, seen)
        if per_ns >= 10:
            assert seen["conflict"] > 0, (per_ns, seen)


def test_c3_literal_match_rate(hostplan, wl):
    """SURVEY.md 8(d) read literally: each of the 100 PodDefaults of a namespace matches a Pod
    with probability 0.3, i.e. 30 applied PodDefaults per Pod on average and 45 at the tail.
    The reference has no limit on how many it applies; nor may the library at these sizes."""
    pods, pds, ranges = wl.admission_workload(240, 100, per_poddefault=True)
    pd_docs = pack(wl, pods, pds)
    outs, st, res = hostplan.poddefault(pd_docs[0], pd_docs[1], ranges)
    assert not (st == 8).any(), "NBRM_E_LIMIT at BASELINE sizes"
    seen = check_poddefault(pods, pds, ranges, outs, st, res)
    applied = [int(r) for r, s_ in zip(res, st) if s_ == 0 and r]
    assert seen["patched"] + seen["conflict"] > 150 and max(applied + [0]) > 32, (seen, max(applied + [0]))
