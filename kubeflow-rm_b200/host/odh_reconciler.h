// odh_reconciler.h — OpenshiftNotebookReconciler: the ODH controller's reconciler in C++
// over the store and the GPU batch engine.
//
// Mirrors components/odh-notebook-controller/controllers/notebook_controller.go:
// Reconcile :150-247, ReconciliationLockIsEnabled :108-115, RemoveReconciliationLock
// :118-146; and the per-object reconcilers it calls: ReconcileAllNetworkPolicies
// (notebook_network.go:43-63, reconcileNetworkPolicy :67-121), ReconcileRoleBindings
// (notebook_rbac.go:144-150), ReconcileOAuthServiceAccount / Service / Secret
// (notebook_oauth.go:77-113, :151-187, :214-250), ReconcileOAuthRoute / ReconcileRoute
// (notebook_route.go:76-139).  As with NotebookReconciler the unit of work is a batch of
// keys: the desired objects of all of them are generated in one GPU pass
// (nbrm_odh_generate), what each reconciler would write given the object it Got is decided
// in a second one (generator 9, nbrm_odhcmp.cuh), and the host applies the writes.
//   * Update conflicts are retried as retry.RetryOnConflict(retry.DefaultRetry, ...) does
//     (notebook_route.go:119, notebook_network.go:101): up to 5 attempts, each re-reading
//     the found object and deciding again.
//   * The lock: while the Notebook carries kubeflow-resource-stopped =
//     odh-notebook-controller-lock, the reconciler waits for the ServiceAccount's image pull
//     secret (3 attempts, 1 s then 5 s apart — `sleep` is the caller's, simulations pass a
//     counter) and then sends the RFC 7386 merge patch
//     {"metadata":{"annotations":{"kubeflow-resource-stopped":null}}} whether or not the
//     secret appeared (the reference discards retry.OnError's result).
// Not built here: CreateNotebookCertConfigMap / UnsetNotebookCertConfig (:254-455; their
// pure part is nbrm_host's ca_bundle functions), checkRoleExists (the caller says whether the
// Role exists through `rbac`).
#pragma once
#include <functional>
#include <string>
#include <vector>

#include "notebook_reconciler.h"

namespace nbrm_host {

struct OdhStats {
  uint64_t keys = 0, reconciled = 0, not_found = 0, creates = 0, updates = 0, conflicts_retried = 0;
  uint64_t locks_removed = 0, pull_secret_waits = 0, errors = 0;
};

class OpenshiftNotebookReconciler {
 public:
  OpenshiftNotebookReconciler(Store* store, nbrm_ctx* ctx, bool rbac, std::function<void(double)> sleep)
      : store_(store), ctx_(ctx), rbac_(rbac), sleep_(std::move(sleep)) {}
  int ReconcileBatch(const std::vector<Request>& reqs, OdhStats* stats, std::vector<std::string>* errors);

 private:
  Store* store_;
  nbrm_ctx* ctx_;
  bool rbac_;
  std::function<void(double)> sleep_;
};

}  // namespace nbrm_host
