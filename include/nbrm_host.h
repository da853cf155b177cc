/* nbrm_host.h — C ABI of libnbrm_host: the C++ host side above libnbrm (an
 * in-memory API object store and the NotebookReconciler that batches keys onto
 * the GPU).  It stands where controller-runtime's client and the reference's
 * NotebookReconciler stand (components/notebook-controller/controllers/
 * notebook_controller.go:75-81, :90-272); see the headers under kubeflow-rm_b200/host. */
#ifndef NBRM_HOST_H_
#define NBRM_HOST_H_
#include <stdint.h>

#include "nbrm.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nbrm_store nbrm_store;
typedef struct nbrm_reconciler nbrm_reconciler;

/* Store: error codes 0 ok, 1 NotFound, 2 AlreadyExists, 3 Conflict, 4 Invalid. */
nbrm_store* nbrm_store_new(void);
void nbrm_store_free(nbrm_store* s);
int nbrm_store_create(nbrm_store* s, const char* json, uint64_t len);
int nbrm_store_update(nbrm_store* s, const char* json, uint64_t len);
int nbrm_store_delete(nbrm_store* s, const char* kind, const char* ns, const char* name);
/* *out is malloc'ed; release with nbrm_store_free_string (NOT nbrm_host_free of nbrm.h, which
 * releases pinned memory of the CUDA runtime: the two libraries are meant to be linked together). */
int nbrm_store_get(nbrm_store* s, const char* kind, const char* ns, const char* name, char** out,
                   uint64_t* len);
/* counters[6] = gets, creates, updates, status_updates, deletes, conflicts */
void nbrm_store_counters(nbrm_store* s, uint64_t* counters, int reset);
uint64_t nbrm_store_size(nbrm_store* s);
void nbrm_store_free_string(char* p);

/* NotebookReconciler.  keys = "namespace/name" entries separated by '\n'.
 * stats[11] = keys, reconciled, not_found, terminating, creates, updates,
 * status_updates, status_unchanged, pod_restarts, errors, events_reissued.  Returns 0, or the
 * number of keys that ended in an error (the reference's `return ctrl.Result{}, err`). */
nbrm_reconciler* nbrm_reconciler_new(nbrm_store* s, nbrm_ctx* ctx, int use_istio, int status_write_if_changed);
void nbrm_reconciler_free(nbrm_reconciler* r);
int nbrm_reconciler_reconcile(nbrm_reconciler* r, const char* keys, uint64_t now_unix, uint64_t* stats);

#ifdef __cplusplus
}
#endif
#endif
