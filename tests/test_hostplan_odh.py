"""CPU: ODH reconcile planners (compiled by g++) against the oracle."""
import json
import os
import random

from parity import check_odh_generate

CTRL_NS = "redhat-ods-applications"


def _seeds(n, seed):
    return random.Random(seed).randbytes(16 * n)


def test_fuzz_corpus(hostplan, wl):
    nbs = [nb for nb in wl.corpus("fuzz", 2000, start=7)]
    rng = random.Random(3)
    docs = [wl.dumps(nb, rng) for nb in nbs]
    seeds = _seeds(len(nbs), 1)
    for flags in (8 | 16, 8 | 16 | 32, 0):
        outs, st = hostplan.odh_generate(docs, seeds, {"CONTROLLER_NAMESPACE": CTRL_NS}, flags)
        assert check_odh_generate(nbs, outs, st, CTRL_NS, seeds, flags) > 6000


def test_reference_fixture(hostplan, wl):
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "odh.json")))
    nbs = [g["oauth"]["notebook"], g["route"]["notebook"], g["network_policies"]["notebook"]] + wl.corpus("c2", 50)
    outs, st = hostplan.odh_generate([wl.dumps(nb) for nb in nbs], None, {"CONTROLLER_NAMESPACE": CTRL_NS}, 8 | 16)
    assert check_odh_generate(nbs, outs, st, CTRL_NS, None, 8 | 16) > 250
    # and directly against the reference's expected structs
    from oracle import odh
    exp = g["oauth"]
    assert odh.compare_notebook_routes(json.loads(outs[2]), exp["expected_route"])
    assert json.loads(outs[3])["metadata"]["annotations"] == exp["expected_service_account"]["metadata"]["annotations"]
    assert json.loads(outs[4])["spec"]["ports"] == exp["expected_service"]["spec"]["ports"]
    assert len(json.loads(outs[5])["stringData"]["cookie_secret"]) == exp["expected_secret_cookie_len"]
    assert odh.compare_notebook_routes(json.loads(outs[7 * 1 + 2]), g["route"]["expected"])
    assert odh.compare_notebook_network_policies(json.loads(outs[7 * 2 + 0]), g["network_policies"]["expected_ctrl"])
    assert odh.compare_notebook_network_policies(json.loads(outs[7 * 2 + 1]), g["network_policies"]["expected_oauth"])
