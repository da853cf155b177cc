#!/usr/bin/env python
"""bench.py — Notebook reconciles/sec on B200 (BASELINE.json metric).

A "step" is one full reconcile pass over one batch of synthetic Notebooks of
BASELINE.json configs[1] (the ODH OAuth Notebook fixture): for every Notebook,
all the pure work the reference does between admission and the end of both
controllers' Reconcile calls (STAGES below; kubeflow-rm_b200/reconcile.py),
against a steady-state cluster (found objects and Pods resident in HBM, as the
reference keeps them in its informer cache).  One Notebook = one reconcile.

  value  reconciles/s with the requests already resident in HBM (kernels only)
  e2e    reconciles/s through the C ABI with HOST buffers: pinned host ->
         device copy of the requests, kernels, device -> host copy of every
         result (mutated Notebooks, the write sets of both controllers,
         statuses), all timed; copies run on their own streams beside the kernels
  --impl reference  the oracle (CPU restatement of the reference; the Go
         reference cannot be built here) on all host cores, same workload

The batch is --batch Notebooks per GPU (default 131 072: 90 MB of requests in,
about 0.9 GB of objects out per step, both larger than the 126 MB L2); the
literal 100-Notebook case of configs[1] is reported as latency in
`config.literal_100`.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Notebook reconciles/sec"
UNIT = "reconciles/s"
ENV = {"USE_ISTIO": "true", "ISTIO_GATEWAY": "kubeflow/kubeflow-gateway", "ISTIO_HOST": "*",
       "CLUSTER_DOMAIN": "cluster.local"}
ENV["CONTROLLER_NAMESPACE"] = "redhat-ods-applications"
NOW = 1793966400   # 2026-11-06T12:00:00Z, the injected clock
WORKLOAD = ("configs[1] odh-notebook-controller: OAuth Notebook fixture (100 per namespace) through webhook, "
            "ODH reconcile and notebook-controller reconcile, steady-state cluster")
STAGES = ["NotebookWebhook.Handle(CREATE): lock + OAuth sidecar",
          "ODH desired objects: NetworkPolicies, Route, ServiceAccount, Service, Secret",
          "ODH write decision vs found objects (reconcileRoute / reconcileNetworkPolicy / create-once)",
          "generateStatefulSet/Service/VirtualService + SetControllerReference",
          "Copy*Fields create-or-update decision vs found objects", "createNotebookStatus"]


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=131072, help="Notebooks per GPU per step")
    ap.add_argument("--ref-sample", type=int, default=8000,
                    help="Notebooks per step of the CPU reference arm")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks and throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        """index: one GPU index, or a comma-separated list (rank 0 samples every GPU of the
        job in one nvidia-smi call: a sampler per rank means N nvidia-smi processes at a
        time, each initialising NVML on all GPUs, beside N ranks that synchronise with
        their GPU six times per step)."""
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self.stop_flag = threading.Event()

    def run(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                     "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                for line in out.strip().splitlines():
                    f = [x.strip() for x in line.split(",")]
                    self.samples.append((float(f[0]), float(f[1])))
                    for nm, v in zip(names, f[2:6]):
                        if v.lower().startswith("active"):
                            self.reasons.add(nm)
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def result(self):
        self.stop_flag.set()
        if self.is_alive():
            self.join(timeout=10)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": sorted(self.reasons)}
        return {"sm_mhz": statistics.median(s[0] for s in self.samples),
                "sm_max_mhz": max(s[1] for s in self.samples), "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def _oracle_chunk(items):
    """Worker: parse the request and its cluster state, run the oracle's full pass,
    serialise every result (the same bytes-in, bytes-out contract as the C ABI)."""
    from oracle import gojson, odh, passes
    now_str = gojson.format_rfc3339(NOW * 10**9)
    n = 0
    for req, found, pod, seed, odh_found in items:
        r = json.loads(req)
        f = [json.loads(x) if x else None for x in found]
        p = json.loads(pod) if pod else None
        of = [json.loads(x) if x else None for x in odh_found]
        mutated, objs, writes, status = passes.full_pass(r, f, f[0], p, ENV, now_str,
                                                         odh.cookie_secret_from_seed(seed), odh_found=of)
        for o in [mutated, status] + [x for x in objs if x is not None] + [x for x in writes if x is not None]:
            n += len(json.dumps(o, separators=(",", ":")))
    return n


def cpu_baseline(items, procs):
    """Oracle timed on host cores over `items`; returns reconciles/s."""
    if procs <= 1:
        t0 = time.perf_counter()
        _oracle_chunk(items)
        return len(items) / (time.perf_counter() - t0)
    import multiprocessing as mp
    chunks = [items[i::procs] for i in range(procs)]
    with mp.get_context("fork").Pool(procs) as pool:
        pool.map(_oracle_chunk, [c[:8] for c in chunks])  # start the workers
        t0 = time.perf_counter()
        pool.map(_oracle_chunk, chunks)
        dt = time.perf_counter() - t0
    return len(items) / dt


def build_workload(wl, pkg, n, start, need_items, need_docs=True):
    """configs[1] x n: OAuth Notebook requests plus the steady-state cluster around
    them (found objects as generated, a running Pod per Notebook).  The found objects
    come from the oracle for the first Notebook and are re-keyed by text substitution
    (every C2 Notebook has the same shape), so building 262 144 of them stays cheap."""
    import random
    from oracle import nbc, odh
    rng = random.Random(wl.SEED + start)
    t_nb = wl.notebook_c2(0)
    mutated, _ = odh.mutate_notebook(t_nb, "CREATE")
    desired = [nbc.generate_statefulset(mutated, ENV), nbc.generate_service(mutated, ENV),
               nbc.generate_virtualservice(mutated, ENV)]
    for d in desired:
        nbc.set_controller_reference(mutated, d)
        d["metadata"]["resourceVersion"] = "7"
    desired[0]["status"] = {"replicas": 1, "readyReplicas": 1}
    pod = {"apiVersion": "v1", "kind": "Pod",
           "metadata": {"name": "nb-000000-0", "namespace": "ns-000", "labels": {"notebook-name": "nb-000000"}},
           "status": {"phase": "Running",
                      "conditions": [{"type": t, "status": "True", "lastProbeTime": None,
                                      "lastTransitionTime": "2026-11-06T11:00:00Z"}
                                     for t in ("Initialized", "Ready", "ContainersReady", "PodScheduled")],
                      "containerStatuses": [{"name": "nb-000000", "ready": True, "restartCount": 0,
                                             "state": {"running": {"startedAt": "2026-11-06T11:00:05Z"}}},
                                            {"name": "oauth-proxy", "ready": True, "restartCount": 0,
                                             "state": {"running": {"startedAt": "2026-11-06T11:00:06Z"}}}]}}
    # what the ODH reconcilers Get: the objects as created, plus what the servers add
    t_odh = []
    for o in odh.reconcile_objects(mutated, ENV["CONTROLLER_NAMESPACE"], rbac=False, cookie_secret="c3RlYWR5"):
        if o is not None:
            o["metadata"]["resourceVersion"] = "7"
            if o["kind"] == "Route":
                o["spec"]["host"] = "nb-000000-ns-000.apps.example.com"
                o["status"] = {"ingress": [{"host": o["spec"]["host"], "routerName": "default"}]}
        t_odh.append(b"" if o is None else wl.dumps(o))
    t_req = wl.dumps(t_nb)
    t_found = [wl.dumps(d) for d in desired]
    t_pod = wl.dumps(pod)
    uid0 = t_nb["metadata"]["uid"].encode()

    def rekey(b, i):
        md = wl._meta(i)
        return (b.replace(b"nb-000000", md["name"].encode()).replace(b"ns-000", md["namespace"].encode())
                .replace(uid0, md["uid"].encode()).replace(b'"resourceVersion":"1000"', b'"resourceVersion":"%d"' % (1000 + i)))
    reqs, found, stss, pods, items, odh_found = [], [], [], [], [], []
    seeds = rng.randbytes(16 * n)
    for k in range(n):
        i = start + k
        r = rekey(t_req, i)
        f = [rekey(x, i) for x in t_found]
        p = rekey(t_pod, i)
        of = [rekey(x, i) for x in t_odh]
        if need_docs:
            odh_found += of
            reqs.append(r)
            found += f
            stss.append(f[0])
            pods.append(p)
        if need_items and k < need_items:
            items.append((r, f, p, seeds[16 * k:16 * k + 16], of))
    return reqs, found, stss, pods, seeds, items, odh_found


def run_reference(args):
    """--impl reference: the oracle on all host cores (rank 0 only)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import __graft_entry__ as graft
    pkg = graft.load_pkg()
    import importlib
    wl = importlib.import_module(graft.PKG_NAME + ".workload")
    cores = os.cpu_count() or 1
    n = args.ref_sample
    items = build_workload(wl, pkg, n, 0, n, need_docs=False)[5]
    for _ in range(min(args.warmup, 1)):
        cpu_baseline(items[: max(n // 10, 1)], cores)
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        cpu_baseline(items, cores)
        times.append(time.perf_counter() - t0)
    dt = sum(times)
    v = n * args.steps / dt
    sample = "%d Notebooks/step (configs[1] fixture + steady-state cluster), JSON in -> oracle full pass -> JSON out, %d processes" % (n, cores)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "stages": STAGES, "batch_notebooks": n},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import __graft_entry__ as graft
    pkg = graft.load_pkg()
    import importlib
    wl = importlib.import_module(graft.PKG_NAME + ".workload")
    rmod = importlib.import_module(graft.PKG_NAME + ".reconcile")

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        os.environ["NCCL_DEBUG"] = "WARN"   # keep NCCL's version banner off stdout: one JSON line only
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    B = args.batch
    # weak scaling: every rank reconciles its own B Notebooks (independent keys, no collective)
    cb_n = 24000   # about 10 s of oracle work on one core
    reqs, found, stss, pods, seeds, items, odh_found = build_workload(wl, pkg, B, rank * B, cb_n if rank == 0 else 0)
    d_req = pkg.Docs(reqs, pinned=True)
    ctx = pkg.Context(local, env=ENV)
    rp = rmod.ReconcilePass(ctx, NOW)
    rp.set_cluster_state(pkg.Docs(found), pkg.Docs(stss), pkg.Docs(pods), odh_found=pkg.Docs(odh_found))
    del found, stss, pods, odh_found

    # ---- value: requests and cluster state resident in HBM, kernels only
    rp.upload_requests(d_req, seeds)
    for _ in range(args.warmup):
        rp.run_resident()
    ctx.sync()
    ctx.profile_enable(True)
    # rank 0 samples the clocks of every GPU of the job; NBRM_BENCH_SAMPLER=all: one per rank
    per_rank = os.environ.get("NBRM_BENCH_SAMPLER", "rank0") == "all"
    sampler = ClockSampler(local if per_rank else ",".join(str(i) for i in range(world)))
    if per_rank or rank == 0:
        sampler.start()
    l0 = ctx.launch_count()
    barrier()
    ctx.timer_start()
    for _ in range(args.steps):
        rp.run_resident()
    ms = ctx.timer_stop()
    barrier()
    launches = ctx.launch_count() - l0
    prof = ctx.profile_read()
    ctx.profile_enable(False)
    ms = max_over_ranks(ms)
    value = world * B * args.steps / (ms / 1e3)

    # ---- e2e: host buffers through the C ABI, copies inside the timed region
    # Every step copies its requests from pinned host memory and reads every result back;
    # as a controller draining its queue would, the copy of step i+1's requests runs on the
    # upload stream while step i is reconciled (ReconcilePass.run_pipelined).
    for out in rp.run_pipelined([(d_req, seeds)] * args.warmup):
        pass
    barrier()
    ctx.timer_start()
    t0 = time.perf_counter()
    for out in rp.run_pipelined([(d_req, seeds)] * args.steps):
        pass
    ms_e2e_dev = ctx.timer_stop()
    ms_e2e = 1e3 * (time.perf_counter() - t0)
    barrier()
    clocks = sampler.result()
    ms_e2e = max_over_ranks(max(ms_e2e, ms_e2e_dev))
    e2e_value = world * B * args.steps / (ms_e2e / 1e3)
    h2d = d_req.nbytes + d_req.off.nbytes + len(seeds)
    d2h = sum(out[k] for k in ("notebooks_bytes", "odh_bytes", "writes_bytes", "status_bytes")) \
        + 8 * (B + 1) + 8 * (7 * B + 1) + 8 * (3 * B + 1) + 8 * (B + 1) + 4 * 2 * (7 + 3 + 1) * B
    assert out["writes_bytes"] == 0 and out["odh_bytes"] == 0, "steady state expected: nothing to create or update"

    # ---- roofline of the dominant kernel (CUDA events inside the value loop, per generator)
    peak, peak_kind = load_peaks()
    names = ["nbc", "odh", "copy_fields", "status", "cull", "mutate", "guard", "poddefault", "profile", "odh_decide"]
    kern = {}
    for g, nm in enumerate(names):
        if prof.runs_gen[g]:
            kern["plan_kernel<%s>" % nm] = (prof.plan_ms_gen[g] / args.steps, prof.plan_bytes_gen[g] / args.steps)
            kern["emit_objects[%s]" % nm] = (prof.emit_ms_gen[g] / args.steps, prof.emit_bytes_gen[g] / args.steps)
    dom = max(kern, key=lambda k: kern[k][0])
    dom_ms, dom_bytes = kern[dom]
    achieved = dom_bytes / (dom_ms / 1e3) / 1e9 if dom_ms > 0 else 0.0
    # DRAM bytes per launch of plan_kernel<copy_fields> from the committed ncu --set full capture
    # (profiles/r01_ncu_full_v5.txt: 924.4 MB read + 24.5 MB written at 65 536 Notebooks)
    traffic = (924.398336e6 + 24.5248e6) / 65536 * B if dom == "plan_kernel<copy_fields>" else None
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_kind": peak_kind,
                "avg_launch_ms": dom_ms, "algorithmic_bytes_per_launch": dom_bytes,
                "all_kernels": {k: {"ms": round(v[0], 4), "GBps": round(v[1] / (v[0] / 1e3) / 1e9, 1) if v[0] else 0}
                                for k, v in kern.items()},
                "kernels_ms_per_step": {"plan": prof.plan_ms / args.steps, "scan": prof.scan_ms / args.steps,
                                        "emit": prof.emit_ms / args.steps}}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # literal configs[1]: 100 Notebooks, latency of one full pass through the C ABI
    small = pkg.Docs(reqs[:100], pinned=True)
    w100 = build_workload(wl, pkg, 100, 0, 0)
    rp.set_cluster_state(*[pkg.Docs(x) for x in w100[1:4]], odh_found=pkg.Docs(w100[6]))
    for _ in range(5):
        rp.run(small, seeds[:1600], copy=False, overlap=True)
    lat = []
    for _ in range(50):
        t0 = time.perf_counter()
        rp.run(small, seeds[:1600], copy=False, overlap=True)
        lat.append(1e3 * (time.perf_counter() - t0))
    lat.sort()

    # CPU baseline: the oracle, one core, bounded sample
    cb = cpu_baseline(items, 1)

    print(json.dumps({
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD + " x %d per GPU" % B,
                   "stages": STAGES, "batch_notebooks": B,
                   "l2": "per step %.0f MB of requests in, %.0f MB walked by the plan kernels (requests, generated "
                         "objects, resident cluster state), %.0f MB of objects written to HBM: larger than L2"
                   % (d_req.nbytes / 1e6, prof.plan_bytes_read / args.steps / 1e6,
                      prof.emit_bytes_written / args.steps / 1e6),
                   "literal_100": {"p50_ms": lat[len(lat) // 2], "p99_ms": lat[-1],
                                   "reconciles_per_s": 100 / (lat[len(lat) // 2] / 1e3)}},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
        "cpu_baseline": {"value": cb, "unit": UNIT, "cores": 1, "kind": "port",
                         "sample": "%d Notebooks of the same workload, JSON in -> oracle full pass -> JSON out" % len(items)},
    }))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
