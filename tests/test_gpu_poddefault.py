"""GPU: PodDefault admission through the C ABI against the oracle (BASELINE config 3 shapes)."""
import pytest

from parity import check_poddefault

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("per_ns", [1, 10, 100])
def test_c3_workload(ctx, pkg, wl, hostplan, per_ns):
    pods, pds, ranges = wl.admission_workload(1000, per_ns)
    a, b = [wl.dumps(p) for p in pods], [wl.dumps(p) for p in pds]
    res = ctx.poddefault_mutate(pkg.Docs(a), pkg.Docs(b), ranges)
    seen = check_poddefault(pods, pds, ranges, res.docs(), res.status, res.result)
    assert seen["patched"] > 50 and seen["none"] > 50, seen
    outs, st, rs = hostplan.poddefault(a, b, ranges)
    assert outs == res.docs() and list(st) == list(res.status) and list(rs) == list(res.result)


def test_no_ranges_and_empty(ctx, pkg, wl):
    pods, pds, _ = wl.admission_workload(40, 5, pods_per_ns=40)
    res = ctx.poddefault_mutate(pkg.Docs([wl.dumps(p) for p in pods]), pkg.Docs([wl.dumps(p) for p in pds]))
    check_poddefault(pods, pds, None, res.docs(), res.status, res.result)
    res = ctx.poddefault_mutate(pkg.Docs([wl.dumps(p) for p in pods]), pkg.Docs([]))
    assert all(d == b"" for d in res.docs()) and int(res.status.sum()) == 0


def test_c3_literal_match_rate(ctx, pkg, wl):
    """1 000 admission requests, 100 PodDefaults per namespace, each matching with probability
    0.3 (SURVEY.md 8(d) read literally): about 30 applied per Pod, none over the fixed room."""
    pods, pds, ranges = wl.admission_workload(1000, 100, per_poddefault=True)
    a, b = [wl.dumps(p) for p in pods], [wl.dumps(p) for p in pds]
    res = ctx.poddefault_mutate(pkg.Docs(a), pkg.Docs(b), ranges)
    assert not (res.status == 8).any(), "NBRM_E_LIMIT at BASELINE sizes"
    seen = check_poddefault(pods, pds, ranges, res.docs(), res.status, res.result)
    applied = [int(r) for r, s in zip(res.result, res.status) if s == 0]
    assert seen["patched"] + seen["conflict"] > 800 and max(applied) > 32, (seen, max(applied))
