"""Times plan_kernel<copy_fields> per kind of pair (StatefulSet / Service / VirtualService
only, and the mix the bench uses) on one GPU:  python profiles/probe_copy.py [N]"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402
import bench  # noqa: E402

pkg = graft.load_pkg()
wl = importlib.import_module(graft.PKG_NAME + ".workload")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
reqs, found, stss, pods, seeds, items, odh_found = bench.build_workload(wl, pkg, N, 0, 0)
ctx = pkg.Context(0, env=bench.ENV)
ctx.upload(0, pkg.Docs(reqs))
ctx.run(5, 0, flags=1)           # mutate
ctx.result_to_slot(0)
ctx.run(0, 0, flags=7 | 8)       # nbc
desired = ctx.download().docs()
out = {"mode": os.environ.get("NBRM_PLAN_MODE", "auto"), "index_found": os.environ.get("PROBE_INDEX", "0")}
for name, sel in (("sts", [0]), ("svc", [1]), ("vs", [2]), ("mix", [0, 1, 2])):
    d = [desired[3 * i + k] for i in range(N) for k in sel]
    f = [found[3 * i + k] for i in range(N) for k in sel]
    ctx.upload(0, pkg.Docs(d))
    ctx.upload(1, pkg.Docs(f))
    if out["index_found"] == "1":
        ctx.slot_index(1)
    for _ in range(2):
        ctx.run(2, 0, 1)
    ctx.sync()
    ctx.profile_enable(True)
    for _ in range(5):
        ctx.run(2, 0, 1)
    p = ctx.profile_read()
    ctx.profile_enable(False)
    res = ctx.download()
    ms = p.plan_ms_gen[2] / 5
    nbytes = sum(len(x) for x in d) + sum(len(x) for x in f)
    out[name] = {"pairs": len(d), "ms": round(ms, 4), "ns_per_pair": round(1e6 * ms / len(d), 1),
                 "GBps": round(nbytes / ms / 1e6, 1), "writes": int(res.off[-1]), "bad_status": int((res.status != 0).sum())}
print(json.dumps(out))
