"""CPU: the ODH write decision (compiled by g++) against the oracle's write_decision."""
import json
import random

from oracle import odh
from parity import check_odh_decide

CTRL_NS = "redhat-ods-applications"


def make_odh_pairs(wl, n, seed):
    """Desired ODH objects from the oracle (7 slots per Notebook, None where the reconcile
    builds nothing), found objects from workload.odh_found_variant."""
    rng = random.Random(seed)
    desired, found = [], []
    for nb in wl.corpus("fuzz", n, start=seed):
        objs = odh.reconcile_objects(nb, CTRL_NS, rbac=True, cookie_secret="c2VjcmV0")
        for o in objs:
            desired.append(o)
            found.append(None if o is None else wl.odh_found_variant(rng, o))
    return desired, found


def pack(wl, desired, found, rng=None):
    d = [b"" if x is None else wl.dumps(x) for x in desired]
    f = [b"" if x is None else wl.dumps(x, rng) for x in found]
    return d, f


def test_odh_decide_fuzz(hostplan, wl):
    desired, found = make_odh_pairs(wl, 800, 61)
    d, f = pack(wl, desired, found, random.Random(9))
    outs, st, res = hostplan.odh_decide(d, f)
    c, u, z = check_odh_decide(desired, found, outs, st, res)
    assert c > 300 and u > 500 and z > 1500, (c, u, z)


def test_odh_decide_edges(hostplan, wl):
    nb = wl.notebook_c2(3)
    objs = odh.reconcile_objects(nb, CTRL_NS, rbac=True, cookie_secret="eA==")
    route = next(o for o in objs if o and o["kind"] == "Route")
    # the ingress controller's host, status and server metadata do not make a Route differ
    f = json.loads(json.dumps(route))
    f["spec"]["host"] = "x.apps"
    f["status"] = {"ingress": []}
    f["metadata"]["resourceVersion"] = "9"
    # pretty-printed and escaped spellings are the same object
    texts = [wl.dumps(f), json.dumps(f, indent=2).encode(), json.dumps(f).replace("edge", "ed\\u0067e").encode()]
    outs, st, res = hostplan.odh_decide([wl.dumps(route)] * 3, texts)
    assert outs == [b""] * 3 and list(st) == [0, 0, 0] and list(res) == [0, 0, 0]
    # a changed port is an update that keeps the found object's own metadata
    g = json.loads(json.dumps(f))
    g["spec"]["port"] = {"targetPort": "other"}
    outs, st, res = hostplan.odh_decide([wl.dumps(route)], [wl.dumps(g)])
    got = json.loads(outs[0])
    assert list(res) == [2] and got["spec"] == route["spec"] and got["metadata"]["resourceVersion"] == "9"
    assert got["status"] == {"ingress": []}
    # a Secret that exists is never touched, whatever it holds
    sec = next(o for o in objs if o and o["kind"] == "Secret")
    outs, st, res = hostplan.odh_decide([wl.dumps(sec)], [b'{"kind":"Secret","stringData":{"cookie_secret":"old"}}'])
    assert outs == [b""] and list(res) == [0]
    # a document that is not an object
    outs, st, res = hostplan.odh_decide([b"[1]"], [b""])
    assert list(st) == [1] and outs == [b""]
