"""Where the e2e milliseconds go: python profiles/e2e_probe.py (one B200)."""
import importlib
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import __graft_entry__ as graft
pkg = graft.load_pkg()
wl = importlib.import_module(graft.PKG_NAME + ".workload")
rmod = importlib.import_module(graft.PKG_NAME + ".reconcile")
B = 131072
START = int(sys.argv[1]) if len(sys.argv) > 1 else 0   # first Notebook index (rank r of bench.py starts at r * B)
reqs, found, stss, pods, seeds, items, odh_found = bench.build_workload(wl, pkg, B, START, 0)
d_req = pkg.Docs(reqs, pinned=True)
ctx = pkg.Context(0, env=bench.ENV)
rp = rmod.ReconcilePass(ctx, bench.NOW)
rp.set_cluster_state(pkg.Docs(found), pkg.Docs(stss), pkg.Docs(pods), odh_found=pkg.Docs(odh_found))
def t(f, n=5):
    for _ in range(3): f()
    ctx.sync(); t0 = time.perf_counter()
    for _ in range(n): f()
    ctx.sync(); return 1e3 * (time.perf_counter() - t0) / n
print("upload only        %.2f ms" % t(lambda: rp.upload_requests(d_req, seeds)))
print("resident, no dl    %.2f ms" % t(lambda: rp.run_resident()))
print("upload+resident    %.2f ms" % t(lambda: (rp.upload_requests(d_req, seeds), rp.run_resident())))
print("resident+dl sync   %.2f ms" % t(lambda: rp.run_resident(download=True, copy=False)))
print("resident+dl overlap %.2f ms" % t(lambda: rp.run_resident(download=True, copy=False, overlap=True)))
print("full e2e overlap   %.2f ms" % t(lambda: rp.run(d_req, seeds, copy=False, overlap=True)))

# is it the kernels that slow down beside the copies, or the copies that wait?
def kernel_ms(f, n=5):
    for _ in range(8): f()
    ctx.sync(); ctx.profile_enable(True); t0 = time.perf_counter()
    for _ in range(n): f()
    ctx.sync(); wall = 1e3 * (time.perf_counter() - t0) / n
    p = ctx.profile_read(); ctx.profile_enable(False)
    return wall, (p.plan_ms + p.scan_ms + p.emit_ms) / n
print("no download:  wall %.2f ms, kernels %.2f ms" % kernel_ms(lambda: rp.run_resident()))
ctx.profile_enable(True)
for _ in range(5): rp.run_resident()
ctx.sync(); p = ctx.profile_read(); ctx.profile_enable(False)
names = ["nbc", "odh", "copy_fields", "status", "cull", "mutate", "guard", "poddefault", "profile", "odh_decide"]
print("plan ms per generator:", " ".join("%s=%.2f" % (nm, p.plan_ms_gen[g] / 5) for g, nm in enumerate(names) if p.runs_gen[g]))
out = rp.run_resident(download=True, copy=True)
print("bytes out:", {k: v for k, v in out.items() if k.endswith("_bytes")}, "status!=0:", {k: int((out[k].status != 0).sum()) for k in ("odh", "writes", "status")})
print("overlapped:   wall %.2f ms, kernels %.2f ms" % kernel_ms(lambda: rp.run_resident(download=True, copy=False, overlap=True)))
def only(names):
    c = ctx
    def f():
        c.run(pkg.G_MUTATE, rmod.SLOT_REQUESTS, flags=pkg.MUT_CREATE)
        c.result_to_slot(rmod.SLOT_NOTEBOOKS)
        if "notebooks" in names: c.download_slot_begin(rmod.SLOT_NOTEBOOKS, 0)
        c.run(pkg.G_ODH, rmod.SLOT_NOTEBOOKS, flags=rp.odh_flags, side=True)
        c.result_to_slot(rmod.SLOT_ODH_DESIRED)
        c.run(pkg.G_ODH_DECIDE, rmod.SLOT_ODH_DESIRED, rmod.SLOT_ODH_FOUND)
        if "odh" in names: c.download_begin(1)
        c.run(pkg.G_NBC, rmod.SLOT_NOTEBOOKS, flags=rp.nbc_flags)
        c.result_to_slot(rmod.SLOT_DESIRED)
        c.run(pkg.G_COPY, rmod.SLOT_DESIRED, rmod.SLOT_FOUND)
        if "writes" in names: c.download_begin(2)
        c.run(pkg.G_STATUS, rmod.SLOT_NOTEBOOKS, rmod.SLOT_STS, rmod.SLOT_PODS, now=bench.NOW)
        if "status" in names: c.download_begin(3)
        for k, nm in enumerate(("notebooks", "odh", "writes", "status")):
            if nm in names: c.download_wait(k, copy=False)
    return f
for names in (["notebooks"], ["odh"], ["writes"], ["status"], ["odh", "writes", "status"]):
    print("lanes %-28s wall %.2f ms, kernels %.2f ms" % ((",".join(names),) + kernel_ms(only(names))))
