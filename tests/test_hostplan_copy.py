"""CPU: the Copy*Fields planner (compiled by g++) against the oracle."""
import json
import random

from oracle import nbc
from parity import check_copy_fields

ENV = {"USE_ISTIO": "true"}


def make_pairs(wl, n, seed):
    """Desired objects from the oracle's generators, found objects from workload.found_variant."""
    rng = random.Random(seed)
    desired, found = [], []
    for nb in wl.corpus("fuzz", n, start=seed):
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
    return desired, found


def pack_pairs(wl, desired, found, rng=None):
    d = [b"" if x is None else wl.dumps(x) for x in desired]
    f = [b"" if x is None else wl.dumps(x, rng) for x in found]
    return d, f


def test_copy_fields_fuzz(hostplan, wl):
    desired, found = make_pairs(wl, 1500, 21)
    d, f = pack_pairs(wl, desired, found, random.Random(4))
    outs, st = hostplan.copy_fields(d, f)
    c, u, z = check_copy_fields(desired, found, outs, st)
    assert c > 300 and u > 1000 and z > 1000, (c, u, z)


def test_copy_fields_edges(hostplan, wl):
    sts = nbc.generate_statefulset(wl.notebook_c1(1), {})
    # escaped and raw spellings of the same string are equal; a real difference is not
    f1 = json.loads(json.dumps(sts))
    f1["metadata"]["uid"] = "x"
    d = [wl.dumps(sts)] * 3
    f = [json.dumps(f1, ensure_ascii=True).encode(), json.dumps(f1, indent=4).encode(),
         json.dumps(f1).replace("busybox", "busybo\\u0078").encode()]
    outs, st = hostplan.copy_fields(d, f)
    assert outs == [b"", b"", b""] and list(st) == [0, 0, 0]
    f2 = json.loads(json.dumps(f1))
    del f2["spec"]["replicas"]
    outs, st = hostplan.copy_fields([wl.dumps(sts)], [wl.dumps(f2)])
    assert list(st) == [5] and outs == [b""]          # reference dereferences a nil *int32
