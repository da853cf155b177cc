"""Documented fixed limits are reported per document, never silent:
checked on the CPU build of the planners here and through the C ABI on the GPU."""
import json

import pytest


def _fragmented_notebook(wl, n_labels):
    nb = wl.notebook_c2(3)
    nb["metadata"]["labels"] = {"k%03d" % i: "v" for i in range(n_labels)}
    return json.dumps(nb, indent=2).encode()      # pretty-printed: members do not merge into runs


def _many_poddefaults(wl, n):
    pod = {"apiVersion": "v1", "kind": "Pod", "metadata": {"name": "p", "namespace": "ns", "labels": {"a": "b"}},
           "spec": {"containers": [{"name": "c", "image": "i"}]}}
    pds = [{"apiVersion": "kubeflow.org/v1alpha1", "kind": "PodDefault",
            "metadata": {"name": "pd-%d" % i, "namespace": "ns", "resourceVersion": "1"},
            "spec": {"selector": {"matchLabels": {"a": "b"}}, "env": [{"name": "E%d" % i, "value": "x"}]}}
           for i in range(n)]
    return wl.dumps(pod), [wl.dumps(p) for p in pds]


def test_limits_cpu_build(hostplan, wl):
    outs, st = hostplan.nbc_generate([_fragmented_notebook(wl, 300), wl.dumps(wl.notebook_c2(4))], {}, 7)
    assert list(st) == [3, 0] and outs[0] == outs[1] == outs[2] == b"" and outs[3] != b""
    outs, st = hostplan.nbc_generate([_fragmented_notebook(wl, 20)], {}, 7)        # still fits
    assert list(st) == [0] and json.loads(outs[0])["spec"]["template"]["metadata"]["labels"]["k019"] == "v"
    pod, pds = _many_poddefaults(wl, 129)
    outs, st, res = hostplan.poddefault([pod], pds)
    assert list(st) == [8] and outs == [b""]                                       # 129 > 128 applied PodDefaults
    pod, pds = _many_poddefaults(wl, 128)
    outs, st, res = hostplan.poddefault([pod], pds)
    assert list(st) == [0] and int(res[0]) == 128
    env = json.loads(outs[0])[1]["value"]["containers"][0]["env"]
    assert [e["name"] for e in env] == ["E%d" % i for i in range(128)]


def _deep_pair(depth):
    """A VirtualService pair whose specs differ only below `depth` levels of nesting."""
    def nest(leaf):
        v = leaf
        for _ in range(depth):
            v = {"k": v}
        return v
    d = {"apiVersion": "networking.istio.io/v1alpha3", "kind": "VirtualService", "metadata": {"name": "a", "namespace": "n"},
         "spec": {"deep": nest(1)}}
    f = {"apiVersion": "networking.istio.io/v1alpha3", "kind": "VirtualService", "metadata": {"name": "a", "namespace": "n"},
         "spec": {"deep": nest(2)}}
    return d, f


def test_compare_depth_limit_is_reported(hostplan, wl):
    """DeepEqual follows 24 levels of nesting; beyond that the pair gets NBRM_E_LIMIT, not a
    silent "differs" (which would be an Update on every reconcile)."""
    d, f = _deep_pair(10)
    outs, st = hostplan.copy_fields([wl.dumps(d)], [wl.dumps(f)])
    assert list(st) == [0] and json.loads(outs[0])["spec"] == d["spec"]            # a real difference: update
    d, f = _deep_pair(40)
    outs, st = hostplan.copy_fields([wl.dumps(d)], [wl.dumps(f)])
    assert list(st) == [8] and outs == [b""]
    outs, st = hostplan.copy_fields([wl.dumps(d)], [wl.dumps(d)])                  # same text: equal at any depth
    assert list(st) == [0] and outs == [b""]


@pytest.mark.gpu
def test_limits_gpu(ctx, pkg, wl):
    ctx.set_env({})
    res = ctx.nbc_generate(pkg.Docs([_fragmented_notebook(wl, 300), wl.dumps(wl.notebook_c2(4))]), 7)
    assert list(res.status) == [3, 0] and res.doc(0) == b"" and res.doc(3) != b""
    pod, pds = _many_poddefaults(wl, 129)
    res = ctx.poddefault_mutate(pkg.Docs([pod]), pkg.Docs(pds))
    assert list(res.status) == [8] and res.doc(0) == b""
    pod, pds = _many_poddefaults(wl, 128)
    res = ctx.poddefault_mutate(pkg.Docs([pod]), pkg.Docs(pds))
    assert list(res.status) == [0] and int(res.result[0]) == 128
    d, f = _deep_pair(40)
    res = ctx.copy_fields(pkg.Docs([wl.dumps(d)]), pkg.Docs([wl.dumps(f)]))
    assert list(res.status) == [8] and res.doc(0) == b""
