// notebook_reconciler.h — NotebookReconciler: the reference's reconciler,
// restated in C++ over the store and the GPU batch engine.
//
// Mirrors components/notebook-controller/controllers/notebook_controller.go:
// type NotebookReconciler :75-81, Reconcile(ctx, req) (ctrl.Result, error)
// :90-272, updateNotebookStatus :274-288.  Same argument meaning (a request is a
// namespace/name key), same outcomes: nil error and no requeue when the Notebook
// is gone (:126-129) or terminating (:135-137); an error (rate-limited requeue in
// controller-runtime) when a client call fails.  The difference is the unit of
// work: ReconcileBatch takes many keys (each at most once, which keeps
// controller-runtime's guarantee that a key is never reconciled concurrently
// with itself) and runs their pure work — desired objects, create-or-update
// decision, status — on the GPU in one pass; Reconcile(req) is a batch of one.
// A key that names an Event takes the re-emission branch (:95-119) with
// nbNameFromInvolvedObject (:666-690): the Event is re-issued on the Notebook (here: an
// Event object created in the store, as record.EventRecorder would).  Not built: the
// Prometheus counters.
#pragma once
#include <string>
#include <vector>

#include "../../include/nbrm.h"
#include "store.h"

namespace nbrm_host {

struct Request {
  std::string ns, name;
};

struct Result {  // ctrl.Result
  bool requeue = false;
  int64_t requeue_after_seconds = 0;
};

struct BatchStats {
  uint64_t keys = 0, reconciled = 0, not_found = 0, terminating = 0;
  uint64_t creates = 0, updates = 0, status_updates = 0, status_unchanged = 0;
  uint64_t pod_restarts = 0, errors = 0, events_reissued = 0;
};

class NotebookReconciler {
 public:
  // use_istio: os.Getenv("USE_ISTIO") == "true" (:213).  status_write_if_changed:
  // false = Status().Update on every reconcile as the reference does (:287);
  // true = skip the write when the mirrored status is byte-identical.
  NotebookReconciler(Store* store, nbrm_ctx* ctx, bool use_istio, bool status_write_if_changed)
      : store_(store), ctx_(ctx), use_istio_(use_istio), write_if_changed_(status_write_if_changed) {}

  // errors[i] is empty when key i reconciled (or was skipped) cleanly, else the error text.
  int ReconcileBatch(const std::vector<Request>& reqs, uint64_t now_unix, BatchStats* stats,
                     std::vector<std::string>* errors);
  Result Reconcile(const Request& req, uint64_t now_unix, std::string* error);

 private:
  Store* store_;
  nbrm_ctx* ctx_;
  bool use_istio_, write_if_changed_;
};

}  // namespace nbrm_host
