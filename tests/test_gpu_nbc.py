"""GPU: notebook-controller generators through the C ABI against the oracle."""
import json
import random

import numpy as np
import pytest

from parity import check_nbc_generate
from test_hostplan_nbc import ENVS

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("envi", range(len(ENVS)))
def test_fuzz_corpus(ctx, pkg, wl, envi):
    env = ENVS[envi]
    ctx.set_env(env)
    nbs = wl.corpus("fuzz", 3000, start=envi * 3000)
    rng = random.Random(envi)
    docs = pkg.Docs([wl.dumps(nb, rng) for nb in nbs])
    for flags in (7, 7 | 8, 1, 2 | 4 | 8 | 16):
        res = ctx.nbc_generate(docs, flags)
        owner = "kubeflow.org/v1" if flags & 16 else "kubeflow.org/v1beta1"
        n = check_nbc_generate(nbs, res.docs(), res.status, env, flags, owner)
        assert n > 2000


def test_reference_fixtures(ctx, pkg, wl):
    env = {"USE_ISTIO": "true"}
    ctx.set_env(env)
    nbs = wl.corpus("c1", 3) + wl.corpus("c2", 100) + wl.corpus("loadtest", 10)
    res = ctx.nbc_generate(pkg.Docs([wl.dumps(nb) for nb in nbs]), 7 | 8)
    assert check_nbc_generate(nbs, res.docs(), res.status, env, 7 | 8) == 3 * len(nbs)


def test_empty_ragged_and_garbage(ctx, pkg, wl):
    ctx.set_env({})
    res = ctx.nbc_generate(pkg.Docs([]), 7)
    assert res.n == 0 and len(res.status) == 0
    res = ctx.nbc_generate(pkg.Docs([b"", b"[]", b"nonsense", b"{}", b'{"metadata":{"name":"x"}}']), 7)
    assert all(d == b"" for d in res.docs())
    assert list(res.status) == [1, 1, 1, 2, 2]
    # one huge document between small ones
    big = wl.notebook_c2(7)
    big["metadata"]["annotations"]["blob"] = "x" * 300000
    nbs = [wl.notebook_c1(1), big, wl.notebook_c1(2)]
    res = ctx.nbc_generate(pkg.Docs([wl.dumps(nb) for nb in nbs]), 7)
    assert check_nbc_generate(nbs, res.docs(), res.status, {}, 7) == 9


def test_matches_cpu_build_of_planner_bytes(ctx, pkg, wl, hostplan):
    """The GPU bytes equal the bytes the same planner source gives on the CPU."""
    ctx.set_env({"USE_ISTIO": "true"})
    nbs = wl.corpus("fuzz", 2000, start=99)
    docs = [wl.dumps(nb) for nb in nbs]
    res = ctx.nbc_generate(pkg.Docs(docs), 7 | 8)
    outs, st = hostplan.nbc_generate(docs, {"USE_ISTIO": "true"}, 7 | 8)
    assert list(st) == list(res.status)
    assert outs == res.docs()


def test_full_size_properties(ctx, pkg, wl):
    """BASELINE-size batch (102 400 Notebooks): every object parses, names line
    up, lengths add up; resident path equals the host-buffer path."""
    ctx.set_env({"USE_ISTIO": "true"})
    n = 102400
    docs = pkg.Docs([wl.dumps(nb) for nb in wl.corpus("c2", n)], pinned=True)
    res = ctx.nbc_generate(docs, 7 | 8)
    assert res.n == 3 * n and int(res.status.sum()) == 0
    assert int(res.off[-1]) == len(res.data)
    ctx.upload(0, docs)
    ctx.nbc_generate_resident(7 | 8)
    res2 = ctx.download()
    assert np.array_equal(res.off, res2.off) and np.array_equal(res.data, res2.data)
    outs = res.docs()
    for i in random.Random(1).sample(range(n), 500):
        for k, kind in enumerate(("StatefulSet", "Service", "VirtualService")):
            t = json.loads(outs[3 * i + k])
            assert t["kind"] == kind
            assert t["metadata"]["ownerReferences"][0]["name"] == "nb-%06d" % i
    # a checksum of checksums over all bytes: sorted input order does not matter
    import zlib
    assert zlib.crc32(res.data.tobytes()) == zlib.crc32(res2.data.tobytes())
