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
"""
from . import capi

SLOT_NOTEBOOKS, SLOT_FOUND, SLOT_DESIRED, SLOT_STS, SLOT_PODS, SLOT_REQUESTS = 0, 1, 2, 3, 4, 5
STAGES = ["NotebookWebhook.Handle(CREATE): lock + OAuth sidecar",
          "ODH desired objects: 2 NetworkPolicies, Route, ServiceAccount, Service, Secret",
          "generateStatefulSet/Service/VirtualService + SetControllerReference",
          "Copy*Fields create-or-update decision vs found objects",
          "createNotebookStatus"]


class ReconcilePass:
    def __init__(self, ctx, now=1793966400):
        self.ctx = ctx
        self.now = now
        self.nbc_flags = capi.GEN_ALL | capi.GEN_OWNER_REF
        self.odh_flags = capi.GEN_OWNER_REF | capi.GEN_OWNER_V1

    def set_cluster_state(self, found, stss, pods):
        """The informer-cache side: found objects (3 per Notebook), StatefulSets and Pods."""
        self.ctx.upload(SLOT_FOUND, found)
        self.ctx.upload(SLOT_STS, stss)
        self.ctx.upload(SLOT_PODS, pods)

    def upload_requests(self, requests, seeds=None):
        self.ctx.upload(SLOT_REQUESTS, requests)
        self.ctx.upload_bytes(seeds)

    def run_resident(self, download=False, copy=True):
        """The four stages on resident inputs.  Returns the downloaded results when asked
        (copy=False: views into the context's pinned buffer, each valid until the next download)."""
        c = self.ctx
        out = {}
        c.run(capi.G_MUTATE, SLOT_REQUESTS, flags=capi.MUT_CREATE)
        c.result_to_slot(SLOT_NOTEBOOKS)
        if download:
            out["notebooks"] = c.download_slot(SLOT_NOTEBOOKS, copy=copy)
            out["notebooks_bytes"] = int(out["notebooks"].off[-1])
        c.run(capi.G_ODH, SLOT_NOTEBOOKS, flags=self.odh_flags, side=True)
        if download:
            out["odh"] = c.download(copy=copy)
            out["odh_bytes"] = int(out["odh"].off[-1])
        c.run(capi.G_NBC, SLOT_NOTEBOOKS, flags=self.nbc_flags)
        c.result_to_slot(SLOT_DESIRED)
        c.run(capi.G_COPY, SLOT_DESIRED, SLOT_FOUND)
        if download:
            out["writes"] = c.download(copy=copy)
            out["writes_bytes"] = int(out["writes"].off[-1])
        c.run(capi.G_STATUS, SLOT_NOTEBOOKS, SLOT_STS, SLOT_PODS, now=self.now)
        if download:
            out["status"] = c.download(copy=copy)
            out["status_bytes"] = int(out["status"].off[-1])
        return out

    def run(self, requests, seeds=None, copy=True):
        """Host buffers in, host results out (the e2e path)."""
        self.upload_requests(requests, seeds)
        return self.run_resident(download=True, copy=copy)
