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

/* r.Patch(ctx, obj, client.RawPatch(...)): json_patch == 0 applies an RFC 7386 merge patch
 * (what RemoveReconciliationLock sends, odh-notebook-controller/controllers/
 * notebook_controller.go:141-145), json_patch != 0 an RFC 6902 JSON patch (what the webhooks
 * answer with).  0 ok, 1 NotFound, 4 the patch does not apply. */
int nbrm_store_patch(nbrm_store* s, const char* kind, const char* ns, const char* name, const char* patch,
                     uint64_t len, int json_patch);

/* OpenshiftNotebookReconciler (odh-notebook-controller/controllers/notebook_controller.go:150-247):
 * NetworkPolicies, Route or the OAuth objects, RoleBinding when rbac, Update conflicts retried
 * as retry.RetryOnConflict does, then RemoveReconciliationLock (:118-146) as a merge patch.
 * stats[9] = keys, reconciled, not_found, creates, updates, conflicts_retried, locks_removed,
 * pull_secret_waits, errors.  *slept_seconds: what the wait for the ServiceAccount's pull
 * secret would have slept (1 s + 5 s per locked Notebook whose secret never came); nothing
 * actually sleeps.  Returns the number of keys that ended in an error. */
typedef struct nbrm_odh_reconciler nbrm_odh_reconciler;
nbrm_odh_reconciler* nbrm_odh_reconciler_new(nbrm_store* s, nbrm_ctx* ctx, int rbac);
void nbrm_odh_reconciler_free(nbrm_odh_reconciler* r);
int nbrm_odh_reconciler_reconcile(nbrm_odh_reconciler* r, const char* keys, uint64_t* stats, double* slept_seconds);

/* ProfileReconciler (profile-controller/controllers/profile_controller.go:105-334): the eight
 * desired objects of every Profile from one GPU pass, then per Profile what the reference does
 * object by object against the store: the Namespace (create, or label update when the `owner`
 * annotation is the Profile's owner, else a `Failed` condition on the Profile and nothing more),
 * AuthorizationPolicy, ServiceAccounts, RoleBindings, ResourceQuota (created, brought back in
 * line, or deleted when the Profile no longer asks for one), PatchDefaultPluginSpec (which
 * writes the Profile every time), the finalizer.  names: Profile names, one per line.
 * workload_identity: the --workload-identity flag ("" = off); default_labels_json: the
 * namespace-labels file as JSON (the context's nbrm_env.namespace_labels_json must hold the
 * same text).  stats[13] = keys, reconciled, not_found, creates, updates, deletes, rejected,
 * profile_writes, finalizers_added, finalizers_removed, plugins_applied, plugins_revoked,
 * errors.  Plugins are C++ callbacks (host/profile_reconciler.h); through this C entry they
 * are counted and do nothing.  Returns the number of Profiles that ended in an error. */
typedef struct nbrm_profile_reconciler nbrm_profile_reconciler;
nbrm_profile_reconciler* nbrm_profile_reconciler_new(nbrm_store* s, nbrm_ctx* ctx, const char* workload_identity,
                                                     const char* default_labels_json);
void nbrm_profile_reconciler_free(nbrm_profile_reconciler* r);
int nbrm_profile_reconciler_reconcile(nbrm_profile_reconciler* r, const char* names, uint64_t* stats);

/* The work queue between watch events and the reconcilers (controller-runtime's workqueue
 * semantics; kubeflow-rm_b200/host/workqueue.h): a key waits at most once, is never handed
 * out while it is being reconciled, comes back after `seconds` (ctrl.Result{RequeueAfter},
 * notebook-controller/controllers/culling_controller.go:143,168) or after the per-key
 * back-off 5 ms * 2^failures capped at 1000 s.  Times are the caller's clock in seconds.
 * nbrm_workqueue_take: up to max ready keys as "ns/name\n" lines (release with
 * nbrm_store_free_string); every key taken must be given back with nbrm_workqueue_done. */
typedef struct nbrm_workqueue nbrm_workqueue;
nbrm_workqueue* nbrm_workqueue_new(void);
void nbrm_workqueue_free(nbrm_workqueue* q);
void nbrm_workqueue_add(nbrm_workqueue* q, const char* ns, const char* name);
void nbrm_workqueue_add_after(nbrm_workqueue* q, const char* ns, const char* name, double now, double seconds);
double nbrm_workqueue_add_rate_limited(nbrm_workqueue* q, const char* ns, const char* name, double now);
void nbrm_workqueue_forget(nbrm_workqueue* q, const char* ns, const char* name);
void nbrm_workqueue_done(nbrm_workqueue* q, const char* ns, const char* name);
uint64_t nbrm_workqueue_len(nbrm_workqueue* q);
double nbrm_workqueue_next_ready(nbrm_workqueue* q);
uint64_t nbrm_workqueue_take(nbrm_workqueue* q, uint64_t max, double now, char** out);

/* The pure part of the trusted-CA bundle path (odh-notebook-controller/controllers/
 * notebook_controller.go): CreateNotebookCertConfigMap :254-357 decides from the bytes of
 * odh-trusted-ca-bundle (ca-bundle.crt, odh-ca-bundle.crt) and kube-root-ca.crt (ca.crt) what
 * goes into workbench-trusted-ca-bundle["ca-bundle.crt"].  A file pointer NULL = the key is
 * absent; *_configmap_found 0 = the Get returned NotFound.  x509_ok: NULL, or three verdicts
 * of the caller's x509.ParseCertificate for the first block of each file in the order above
 * (1 parses, 0 does not, -1 no verdict: the DER's outer structure is checked instead).
 * Returns 1 with *out = the bundle (malloc'ed, nbrm_store_free_string), 0 when the reference
 * writes no ConfigMap.  The reads and the Create / Update stay with the caller. */
int nbrm_ca_bundle(const char* ca_bundle_crt, const char* odh_ca_bundle_crt, int odh_configmap_found,
                   const char* ca_crt, int kube_configmap_found, const int8_t* x509_ok, char** out,
                   uint64_t* len);
/* CERTIFICATE blocks in a bundle, as the reference's checkCertConfigMap counts them
 * (notebook_controller_test.go:1134-1159). */
uint64_t nbrm_pem_count_certificates(const char* bundle, uint64_t len);
/* UnsetNotebookCertConfig :390-455 on a Notebook's JSON: the five CA environment variables and
 * the trusted-ca mount leave the notebook's own container, the workbench-trusted-ca-bundle
 * volume leaves the pod.  1: *out is the Notebook to patch to; 0: nothing to change; -1: the
 * reference panics (its delete-while-ranging loop :411-415 slices past the shortened list
 * when a variable is listed twice and the second one is last); -2: not a Notebook object. */
int nbrm_unset_notebook_cert_config(const char* notebook, uint64_t nlen, char** out, uint64_t* len);

/* getBindingName (access-management/kfam/bindings.go:61-78): the name of the RoleBinding kfam
 * creates for a contributor, e.g. ("User", "lalith.vaka@zq.msds.kp.org", "ClusterRole", "edit")
 * -> "user-lalith-vaka-zq-msds-kp-org-clusterrole-edit".  malloc'ed: nbrm_store_free_string. */
char* nbrm_kfam_binding_name(const char* user_kind, const char* user_name, const char* roleref_kind,
                             const char* roleref_name);

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
