"""CPU: the planner source (compiled by g++) against the oracle on the fuzz
corpus.  This checks the walker logic without a GPU; the GPU parity tests
(test_gpu_nbc.py) run the same comparison through the C ABI."""
import pytest

from parity import check_nbc_generate

ENVS = [
    {},
    {"USE_ISTIO": "true", "ISTIO_GATEWAY": "kubeflow/kubeflow-gateway", "ISTIO_HOST": "*",
     "CLUSTER_DOMAIN": "cluster.local"},
    {"ISTIO_HOST": "nb.example.com", "ISTIO_GATEWAY": "istio-system/gw\"q", "CLUSTER_DOMAIN": "",
     "ADD_FSGROUP": "false"},
    {"ADD_FSGROUP": "true", "CLUSTER_DOMAIN": "c.internal"},
]


@pytest.mark.parametrize("envi", range(len(ENVS)))
def test_fuzz_corpus(hostplan, wl, envi):
    import random
    env = ENVS[envi]
    nbs = wl.corpus("fuzz", 1500, start=envi * 1500)
    rng = random.Random(envi)
    docs = [wl.dumps(nb, rng) for nb in nbs]
    for flags in (7, 7 | 8, 1, 2 | 4 | 8 | 16):
        outs, st = hostplan.nbc_generate(docs, env, flags)
        owner = "kubeflow.org/v1" if flags & 16 else "kubeflow.org/v1beta1"
        n = check_nbc_generate(nbs, outs, st, env, flags, owner)
        assert n > 1000


def test_reference_fixtures(hostplan, wl):
    nbs = wl.corpus("c1", 3) + wl.corpus("c2", 100) + wl.corpus("loadtest", 10)
    docs = [wl.dumps(nb) for nb in nbs]
    outs, st = hostplan.nbc_generate(docs, {"USE_ISTIO": "true"}, 7 | 8)
    assert check_nbc_generate(nbs, outs, st, {"USE_ISTIO": "true"}, 7 | 8) == 3 * len(nbs)


def test_empty_and_garbage(hostplan):
    outs, st = hostplan.nbc_generate([], {}, 7)
    assert outs == [] and len(st) == 0
    outs, st = hostplan.nbc_generate([b"", b"[]", b"nonsense", b"{}", b'{"metadata":{"name":"x"}}'], {}, 7)
    assert all(o == b"" for o in outs)
    assert list(st) == [1, 1, 1, 2, 2]
