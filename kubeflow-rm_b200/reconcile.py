"""Host-side composition of one full reconcile pass over a batch of Notebooks.

Mirrors, for BASELINE.json configs[1] (ODH Notebooks with OAuth proxy and
NetworkPolicies), what happens to one Notebook between its admission and the end
of both controllers' Reconcile calls, minus every client round trip:

  1. NotebookWebhook.Handle            odh .../notebook_webhook.go:234-302   (CREATE)
  2. OpenshiftNotebookReconciler       odh .../notebook_controller.go:188-232 desired objects
  3. NotebookReconciler.Reconcile      nbc .../notebook_controller.go:139-218 desired objects
                                       + create-or-update decision against the found objects
  4. updateNotebookStatus              nbc .../notebook_controller.go:274-349

Slot layout on the context (all device-resident between passes):
  5 requests (the Notebooks as submitted)     0 mutated Notebooks (output of 1)
  1 found StatefulSet/Service/VirtualService  2 desired objects (scratch of 3)
  3 found StatefulSets (for status)           4 Pods "<name>-0"
  6 found ODH objects, 7 per Notebook (optional)   7 desired ODH objects (scratch of 2)
"""
from . import capi

SLOT_NOTEBOOKS, SLOT_FOUND, SLOT_DESIRED, SLOT_STS, SLOT_PODS, SLOT_REQUESTS = 0, 1, 2, 3, 4, 5
SLOT_ODH_FOUND, SLOT_ODH_DESIRED = 6, 7
SLOT_REQUESTS_B = 8   # second request buffer of run_pipelined
STAGES = ["NotebookWebhook.Handle(CREATE): lock + OAuth sidecar",
          "ODH desired objects: 2 NetworkPolicies, Route, ServiceAccount, Service, Secret",
          "ODH write decision vs found ODH objects (when the cluster state has them)",
          "generateStatefulSet/Service/VirtualService + SetControllerReference",
          "Copy*Fields create-or-update decision vs found objects",
          "createNotebookStatus"]


class ReconcilePass:
    def __init__(self, ctx, now=1793966400):
        self.ctx = ctx
        self.now = now
        self.nbc_flags = capi.GEN_ALL | capi.GEN_OWNER_REF
        self.odh_flags = capi.GEN_OWNER_REF | capi.GEN_OWNER_V1
        self.odh_decide = False

    def set_cluster_state(self, found, stss, pods, odh_found=None, index=True):
        """The informer-cache side: found objects (3 per Notebook), StatefulSets and Pods;
        optionally the found ODH objects (7 per Notebook, empty = NotFound): with them the
        "odh" result is what the ODH reconcilers write (create / update / nothing) instead
        of every desired object."""
        self.ctx.upload(SLOT_FOUND, found)
        self.ctx.upload(SLOT_STS, stss)
        self.ctx.upload(SLOT_PODS, pods)
        self.odh_decide = odh_found is not None
        if self.odh_decide:
            self.ctx.upload(SLOT_ODH_FOUND, odh_found)
        if index:
            # the informer-cache analogue: resident objects are indexed once, when they arrive
            for slot in (SLOT_FOUND, SLOT_STS, SLOT_PODS) + ((SLOT_ODH_FOUND,) if self.odh_decide else ()):
                self.ctx.slot_index(slot)

    def upload_requests(self, requests, seeds=None):
        self.ctx.upload(SLOT_REQUESTS, requests)
        self.ctx.upload_bytes(seeds)

    def run_resident(self, download=False, copy=True, overlap=False, request_slot=SLOT_REQUESTS):
        """The four stages on resident inputs.  Returns the downloaded results when asked
        (copy=False: views into the context's pinned buffers, valid until the next pass).
        overlap=True sends each finished result to the host on the copy stream while the
        next stage runs (nbrm_download_begin / _wait, one lane per result); the results are
        the same bytes either way."""
        c = self.ctx
        out = {}
        lanes = {}

        def fetch(name, lane, slot=None):
            if not download:
                return
            if overlap:
                if slot is None:
                    c.download_begin(lane)
                else:
                    c.download_slot_begin(slot, lane)
                lanes[name] = lane
            else:
                out[name] = c.download(copy=copy) if slot is None else c.download_slot(slot, copy=copy)
                # with copy=False this view is only valid until the next synchronous download
                out[name + "_bytes"] = int(out[name].off[-1])

        c.run(capi.G_MUTATE, request_slot, flags=capi.MUT_CREATE)
        c.result_to_slot(SLOT_NOTEBOOKS)
        fetch("notebooks", 0, SLOT_NOTEBOOKS)
        c.run(capi.G_ODH, SLOT_NOTEBOOKS, flags=self.odh_flags, side=True)
        if self.odh_decide:
            c.result_to_slot(SLOT_ODH_DESIRED)
            c.run(capi.G_ODH_DECIDE, SLOT_ODH_DESIRED, SLOT_ODH_FOUND)
        fetch("odh", 1)
        # the status depends on the mutated Notebook, its StatefulSet and Pod only; it runs
        # before the compare so that its download hides under it (the last result, the
        # write set, is empty in a steady state)
        c.run(capi.G_STATUS, SLOT_NOTEBOOKS, SLOT_STS, SLOT_PODS, now=self.now)
        fetch("status", 3)
        c.run(capi.G_NBC, SLOT_NOTEBOOKS, flags=self.nbc_flags)
        c.result_to_slot(SLOT_DESIRED)
        c.run(capi.G_COPY, SLOT_DESIRED, SLOT_FOUND)
        fetch("writes", 2)
        for name, lane in lanes.items():
            out[name] = c.download_wait(lane, copy=copy)
            out[name + "_bytes"] = int(out[name].off[-1])
        return out

    def run(self, requests, seeds=None, copy=True, overlap=False):
        """Host buffers in, host results out (the e2e path)."""
        self.upload_requests(requests, seeds)
        return self.run_resident(download=True, copy=copy, overlap=overlap)

    def run_pipelined(self, batches, copy=False):
        """batches: [(pinned request Docs, seeds), ...].  Yields each batch's results (host
        buffers in, host results out, like run); the requests of batch i+1 are copied to the
        device on the upload stream while batch i is being reconciled (nbrm_upload_async on
        two alternating slots) and results leave on the download lanes.  With copy=False a
        result is valid until the next batch's results are produced."""
        slots = (SLOT_REQUESTS, SLOT_REQUESTS_B)
        if not batches:
            return
        self.ctx.upload_async(slots[0], batches[0][0])
        for i, (_, seeds) in enumerate(batches):
            if i + 1 < len(batches):
                self.ctx.upload_async(slots[(i + 1) & 1], batches[i + 1][0])
            self.ctx.upload_bytes(seeds)
            yield self.run_resident(download=True, copy=copy, overlap=True, request_slot=slots[i & 1])
