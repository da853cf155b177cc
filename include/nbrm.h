/* nbrm.h — C ABI of libnbrm, the B200 batch engine for the kubeflow-rm hot path.
 *
 * The reference (rhoai-rhtap/kubeflow-rm) is Go and exposes no in-process
 * plugin or FFI interface: its drop-in surface is the CRDs and
 * the webhooks.  This ABI is the seam a Go maintainer would bind with cgo (see
 * INTEGRATION.md) to hand the pure, per-object work of a reconcile or
 * admission pass to the GPU in batches.  Each entry point names the reference
 * function(s) it replaces.  Paths are relative to components/ of the
 * reference.
 *
 * Conventions
 *   - A batch is n UTF-8 JSON documents packed back to back in `data`, with
 *     n+1 byte offsets in `off` (document i is data[off[i] .. off[i+1])).
 *   - All pointers in nbrm_docs are HOST pointers.  Pinned memory
 *     (nbrm_host_alloc) makes the copies asynchronous; pageable memory works.
 *   - Results point into context-owned pinned host memory and stay valid until
 *     the next call on the same context.
 *   - status[i] != 0 means input i was rejected the way the reference would
 *     reject or crash on it (NBRM_E_*); its output objects have length 0.
 *   - Return value: 0 ok; 1 ok but at least one status is non-zero; <0 the
 *     call failed (nbrm_last_error says why).  There is no CPU fallback: every
 *     entry point fails when no CUDA device is usable.
 *   - A context is not thread-safe; use one per worker thread.  Calls on one
 *     context run in order on its own CUDA stream.
 *   - Limits: one batch holds < 4 GiB of input and < 4 GiB of output; one
 *     document < 256 MiB.
 */
#ifndef NBRM_H_
#define NBRM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nbrm_ctx nbrm_ctx;

/* Per-document status codes. */
enum {
  NBRM_OK = 0,
  NBRM_E_PARSE = 1,         /* not a JSON object / required object missing */
  NBRM_E_NO_CONTAINERS = 2, /* reference panics: Containers[0], notebook_controller.go:455,489 */
  NBRM_E_SEGS = 3,          /* object needs more template segments than the plan holds */
  NBRM_E_PORTS_EMPTY = 4,   /* reference panics: Ports[0] on "ports":[], notebook_controller.go:491 */
  NBRM_E_NIL_REPLICAS = 5,  /* reference panics: *Spec.Replicas, common/reconcilehelper/util.go:123 */
  NBRM_E_DENIED = 6,        /* admission.Denied, odh-notebook-controller/controllers/notebook_webhook.go:274 */
  NBRM_E_CONFLICT = 7,      /* PodDefault merge conflict, admission-webhook/main.go:671-679 */
  NBRM_E_LIMIT = 8,         /* a documented fixed limit of this library was exceeded */
  NBRM_E_SELECTOR = 9,      /* label selector conversion failed, admission-webhook/main.go:77-79 */
  NBRM_E_ENV = 10           /* a required environment variable is not set (getEnvVariable,
                             * tensorboard-controller/controllers/tensorboard_controller.go:535-541) or
                             * RWO_PVC_SCHEDULING has a value rwoPVCScheduling :471-482 rejects */
};

/* The process environment and flags the reference reads while generating
 * objects.  NULL means "unset" (os.LookupEnv reports absent).
 *   notebook-controller/controllers/notebook_controller.go:213 (USE_ISTIO),
 *   :475 (ADD_FSGROUP), :537 (CLUSTER_DOMAIN), :548 (ISTIO_HOST), :557 (ISTIO_GATEWAY);
 *   odh-notebook-controller/main.go:63-72 (controller namespace), :80 (--oauth-proxy-image). */
typedef struct nbrm_env {
  const char* use_istio;
  const char* istio_gateway;
  const char* istio_host;
  const char* cluster_domain;
  const char* add_fsgroup;
  const char* controller_namespace;
  const char* oauth_proxy_image;
  /* profile-controller: --userid-header / --userid-prefix (profile-controller/main.go:75-76),
   * NOTEBOOK_CONTROLLER_PRINCIPAL, ISTIO_INGRESS_GATEWAY_PRINCIPAL, KFP_UI_PRINCIPAL
   * (controllers/profile_controller.go:420-430), and the namespace-labels file
   * (--namespace-labels-path, :775-790) given as a JSON object of strings. */
  const char* userid_header;
  const char* userid_prefix;
  const char* notebook_controller_principal;
  const char* istio_ingress_gateway_principal;
  const char* kfp_ui_principal;
  const char* namespace_labels_json;
  /* tensorboard-controller: TENSORBOARD_IMAGE (controllers/tensorboard_controller.go:172),
   * RWO_PVC_SCHEDULING (:471-482).  It reads ISTIO_GATEWAY / ISTIO_HOST (:325,:330) too, and
   * takes them as set, with no default: for nbrm_tensorboard_generate NULL means "not set". */
  const char* tensorboard_image;
  const char* rwo_pvc_scheduling;
  /* pvcviewer-controller webhook: the PodSpec of the defaults file DEFAULT_POD_SPEC_PATH names
   * (api/v1alpha1/pvcviewer_webhook.go:78-91), as JSON; NULL or "" = the variable is not set. */
  const char* pvcviewer_default_podspec_json;
} nbrm_env;

typedef struct nbrm_docs {
  const char* data;
  const uint64_t* off; /* n + 1 entries */
  uint64_t n;
} nbrm_docs;

typedef struct nbrm_out {
  const char* data;
  const uint64_t* off;    /* n + 1 entries */
  uint64_t n;             /* number of output objects */
  const uint32_t* status; /* one per INPUT document (tuple) */
  uint64_t n_status;
  const uint32_t* result; /* one per input tuple; meaning per entry point, 0 when unused */
} nbrm_out;

/* Accumulated device time per kernel class since nbrm_profile_enable(ctx, 1),
 * measured with CUDA events on the context's stream. */
#define NBRM_GENERATORS 32 /* size of the per-generator arrays below (a power of two); ids 0..18 are in use */
typedef struct nbrm_profile {
  double plan_ms, scan_ms, emit_ms;
  uint64_t plan_launches, scan_launches, emit_launches;
  uint64_t emit_bytes_read, emit_bytes_written; /* algorithmic bytes of the emit kernel */
  uint64_t plan_bytes_read;                     /* input bytes the plan kernels walked */
  /* the same per generator id (see nbrm_run): device ms, algorithmic bytes (plan: input
   * bytes read; emit: output bytes read + written), number of runs */
  double plan_ms_gen[NBRM_GENERATORS], emit_ms_gen[NBRM_GENERATORS];
  uint64_t plan_bytes_gen[NBRM_GENERATORS], emit_bytes_gen[NBRM_GENERATORS], runs_gen[NBRM_GENERATORS];
} nbrm_profile;

/* ---- context ---------------------------------------------------------- */
int nbrm_ctx_create(int device, nbrm_ctx** ctx);
void nbrm_ctx_destroy(nbrm_ctx* ctx);
const char* nbrm_last_error(const nbrm_ctx* ctx);
int nbrm_ctx_set_env(nbrm_ctx* ctx, const nbrm_env* env);
/* Options.  NBRM_OPT_GROUP_BY_LENGTH (default 0): for batches of more than 4096 tuples, hand the
 * tuples to the threads of a plan kernel in the order of their first document's length, so that
 * documents of one shape share warps; results do not move.  Worth it on batches that mix many
 * shapes (the objects of a real cluster), costs a sort of n lengths per generator on uniform
 * ones. */
enum { NBRM_OPT_GROUP_BY_LENGTH = 1 };
int nbrm_ctx_set_option(nbrm_ctx* ctx, int option, int value);
int nbrm_sync(nbrm_ctx* ctx);

/* Pinned host memory for inputs (optional). */
void* nbrm_host_alloc(size_t bytes);
void nbrm_host_free(void* p);

/* Device timing of whatever runs on the context's stream between the calls. */
int nbrm_timer_start(nbrm_ctx* ctx);
int nbrm_timer_stop(nbrm_ctx* ctx, float* ms);
int nbrm_profile_enable(nbrm_ctx* ctx, int on);
int nbrm_profile_read(nbrm_ctx* ctx, nbrm_profile* out);
/* Number of kernels this library has launched on the context so far. */
uint64_t nbrm_launch_count(const nbrm_ctx* ctx);

/* ---- resident inputs -------------------------------------------------- */
/* A context holds NBRM_SLOTS input slots.  The named entry points below use
 * slots 0..2 as they document; nbrm_run lets a host compose generators over any
 * slots (e.g. keep the found-object cache and the Pods resident in slots of
 * their own while requests stream through another). */
#define NBRM_SLOTS 12
/* Copy a batch into an input slot of the context; it stays in HBM until the
 * slot is overwritten. */
int nbrm_upload(nbrm_ctx* ctx, int slot, const nbrm_docs* docs);
/* Copy a raw side input (meaning per entry point, e.g. OAuth cookie seeds) to
 * the device; bytes == 0 clears it. */
int nbrm_upload_bytes(nbrm_ctx* ctx, const void* host, uint64_t bytes);
/* Copy the result of the last *_resident call to host memory. */
int nbrm_download(nbrm_ctx* ctx, nbrm_out* out);

/* ---- notebook-controller generators ------------------------------------
 * Replaces generateStatefulSet (notebook-controller/controllers/
 * notebook_controller.go:408-484, with setPrefixEnvVar :392-406),
 * generateService (:486-513), generateVirtualService (:519-619) and, with
 * NBRM_GEN_OWNER_REF, the ctrl.SetControllerReference calls at :141,:183,:628.
 * Input slot 0: Notebook CRs.  Output: 3 objects per Notebook, in the order
 * StatefulSet, Service, VirtualService; an object not asked for has length 0. */
enum {
  NBRM_GEN_STS = 1,
  NBRM_GEN_SVC = 2,
  NBRM_GEN_VS = 4,
  NBRM_GEN_OWNER_REF = 8,
  NBRM_GEN_OWNER_V1 = 16 /* ownerReference apiVersion kubeflow.org/v1 (ODH) instead of v1beta1 */
};
int nbrm_nbc_generate(nbrm_ctx* ctx, const nbrm_docs* notebooks, uint32_t flags, nbrm_out* out);
int nbrm_nbc_generate_resident(nbrm_ctx* ctx, uint32_t flags);

/* ---- create-or-update decision ---------------------------------------------
 * Replaces reconcilehelper.CopyStatefulSetFields / CopyServiceFields /
 * CopyVirtualService (common/reconcilehelper/util.go:107-134, :166-195,
 * :199-219), the template-label pre-copy of notebook_controller.go:165-169 and
 * the NotFound -> Create / changed -> Update decision around them
 * (notebook_controller.go:147-179, :189-210, :636-657).
 * Input slot 0: desired objects; slot 1: found objects, pair i = document i of
 * both, an empty found document = NotFound.  The object kind is read from the
 * desired object's "kind".  Output, one object per pair: the desired object
 * (the caller Creates it), or the found object with the owned fields copied
 * over (the caller Updates it), or length 0 (nothing to do).  status
 * NBRM_E_NIL_REPLICAS marks the pair the reference would crash on. */
int nbrm_copy_fields(nbrm_ctx* ctx, const nbrm_docs* desired, const nbrm_docs* found, nbrm_out* out);
int nbrm_copy_fields_resident(nbrm_ctx* ctx);
/* Move the last result into an input slot without copying it (the slot's old
 * contents are dropped).  Lets one generator's output feed the next on the device. */
int nbrm_result_to_slot(nbrm_ctx* ctx, int slot);

/* One notebook-controller reconcile pass for a batch, on the device end to end:
 * nbrm_nbc_generate, then nbrm_copy_fields against the found objects
 * (NotebookReconciler.Reconcile, notebook_controller.go:139-218, minus the
 * client calls).  Slot 0: Notebooks; slot 1: found objects, 3 per Notebook in
 * the order StatefulSet, Service, VirtualService (empty = NotFound); slot 2 is
 * used for the desired objects.  Output: 3 objects per Notebook as
 * nbrm_copy_fields gives them; in steady state every length is 0 and only the
 * offsets travel back.  The status array then has one entry per object pair; a
 * Notebook the generators rejected (NBRM_E_NO_CONTAINERS, NBRM_E_PORTS_EMPTY,
 * NBRM_E_PARSE, NBRM_E_SEGS: where the reference panics or errors before it
 * writes anything) carries that code on all three of its pairs, and the caller
 * must not go on to the status update for it. */
int nbrm_nbc_reconcile(nbrm_ctx* ctx, const nbrm_docs* notebooks, const nbrm_docs* found,
                       uint32_t flags, nbrm_out* out);
int nbrm_nbc_reconcile_resident(nbrm_ctx* ctx, uint32_t flags);

/* ---- Notebook status mirroring --------------------------------------------
 * Replaces createNotebookStatus and PodCondToNotebookCond (notebook-controller/
 * controllers/notebook_controller.go:290-349, :351-390).  Slot 0: Notebooks,
 * slot 1: their StatefulSets, slot 2: their Pods "<name>-0"; an empty document
 * = not found.  Output: one NotebookStatus object per Notebook (what the
 * reference assigns to nb.Status before Status().Update, :286-287).
 * now_unix stands for metav1.Now() (:377,:386), in Unix seconds. */
int nbrm_notebook_status(nbrm_ctx* ctx, const nbrm_docs* notebooks, const nbrm_docs* statefulsets,
                         const nbrm_docs* pods, uint64_t now_unix, nbrm_out* out);
int nbrm_notebook_status_resident(nbrm_ctx* ctx, uint64_t now_unix);

/* ---- idle culling ---------------------------------------------------------
 * Replaces the pure part of CullingReconciler.Reconcile (notebook-controller/
 * controllers/culling_controller.go:85-169) and its helpers
 * cullingCheckPeriodHasPassed :173, notebookIsIdle :186, allKernelsAreIdle :281,
 * getNotebookRecentTime :295, compareAnnotationTimeToResource :337,
 * updateTimestampFrom{Kernels,Terminals}Activity :357,:390,
 * updateLastCullingCheckTimestampAnnotation :416, annotationsExist :426,
 * initializeAnnotations :435, removeAnnotations :444, setStopAnnotation :461,
 * StopAnnotationIsSet :480.  The HTTP GETs to the notebook server stay with
 * the caller: slot 1 / slot 2 hold the bodies of /api/kernels and
 * /api/terminals per Notebook (empty document = the GET failed), pod_exists one
 * byte per Notebook (NULL = every Pod exists).  cull_idle_minutes and
 * check_period_minutes are CULL_IDLE_TIME and IDLENESS_CHECK_PERIOD (:511-544).
 * Output: the Notebook to Update, length 0 when the pass changes nothing;
 * result[i] = outcome | updates << 8 with outcome 1 culling annotations
 * removed because the Notebook is stopped, 2 removed because the Pod is gone,
 * 3 check period not over (requeue), 4 checked and still active, 5 culled
 * (stop annotation set); updates = number of r.Update calls the reference
 * makes on that path. */
int nbrm_culling_pass(nbrm_ctx* ctx, const nbrm_docs* notebooks, const nbrm_docs* kernels,
                      const nbrm_docs* terminals, const uint8_t* pod_exists, uint64_t now_unix,
                      int32_t cull_idle_minutes, int32_t check_period_minutes, nbrm_out* out);
int nbrm_culling_pass_resident(nbrm_ctx* ctx, uint64_t now_unix, int32_t cull_idle_minutes,
                               int32_t check_period_minutes);

/* ---- ODH reconcile generators -------------------------------------------
 * Replaces, for a batch of Notebooks, the desired-object constructors that
 * OpenshiftNotebookReconciler.Reconcile calls (odh-notebook-controller/
 * controllers/notebook_controller.go:188-232):
 *   NewNotebookNetworkPolicy  notebook_network.go:131-173
 *   NewOAuthNetworkPolicy     notebook_network.go:176-210
 *   NewNotebookRoute          notebook_route.go:34-62
 *   NewNotebookOAuthRoute     notebook_oauth.go:253-259
 *   NewNotebookServiceAccount notebook_oauth.go:49-64
 *   NewNotebookOAuthService   notebook_oauth.go:116-140
 *   NewNotebookOAuthSecret    notebook_oauth.go:190-210
 *   NewRoleBinding            notebook_rbac.go:36-58 (as called at :144-153)
 * and the decisions around them (OAuthInjectionIsEnabled :86, ServiceMeshIsEnabled
 * :97).  Output: 7 objects per Notebook in the order ctrl NetworkPolicy, OAuth
 * NetworkPolicy, Route (plain or OAuth), ServiceAccount, Service, Secret,
 * RoleBinding; an object the reference would not build for that Notebook has
 * length 0.  NBRM_GEN_OWNER_REF | NBRM_GEN_OWNER_V1 add the ownerReference the
 * Reconcile* helpers set before Create.  NBRM_ODH_RBAC stands for
 * SET_PIPELINE_RBAC=="true" with the Role present (notebook_rbac.go:61,144).
 * seeds16: 16 bytes per Notebook from the caller's CSPRNG (the reference draws
 * them with crypto/rand, notebook_oauth.go:192-193); the Secret's cookie is
 * base64(base64(seed)).  NULL seeds give the all-zero seed: tests only. */
enum { NBRM_ODH_RBAC = 32 };
int nbrm_odh_generate(nbrm_ctx* ctx, const nbrm_docs* notebooks, const uint8_t* seeds16,
                      uint32_t flags, nbrm_out* out);
int nbrm_odh_generate_resident(nbrm_ctx* ctx, uint32_t flags);

/* ---- ODH Notebook mutating webhook -------------------------------------------
 * Replaces the object transformation of NotebookWebhook.Handle
 * (odh-notebook-controller/controllers/notebook_webhook.go:234-302):
 * InjectReconciliationLock :63-72 (CREATE only), InjectCertConfig :421-535 (with
 * NBRM_MUT_CA_BUNDLE: the caller found the ConfigMaps CheckAndMountCACertBundle
 * :373-419 reads), InjectOAuthProxy :76-231 when the inject-oauth annotation
 * parses true, admission.Denied (status NBRM_E_DENIED) when service mesh is on
 * too (:272-275), and for UPDATE (old_notebooks != NULL)
 * maybeRestartRunningNotebook :314-370 with the update-pending annotation
 * :288-293.  The OAuth proxy image is nbrm_env.oauth_proxy_image.
 * SetContainerImageFromRegistry (:541-640) needs a live ImageStream list and
 * stays with the caller.  Output: the mutated Notebook per request (the caller
 * diffs it against the request object as admission.PatchResponseFromRaw does).
 * result[i] = 1 when the webhook's own pod-template changes were held back
 * (UpdatesPending).  The update-pending annotation then carries a fixed text:
 * the reference's go-cmp diff string is not reproducible outside Go.
 * Slots used: 1 (request objects), 2 (old objects), 0 (intermediate). */
enum { NBRM_MUT_CREATE = 1, NBRM_MUT_CA_BUNDLE = 2 };
int nbrm_odh_mutate(nbrm_ctx* ctx, const nbrm_docs* notebooks, const nbrm_docs* old_notebooks,
                    uint32_t flags, nbrm_out* out);
int nbrm_odh_mutate_resident(nbrm_ctx* ctx, uint32_t flags, int is_update);

/* ---- PodDefault admission ---------------------------------------------------
 * Replaces, for a batch of Pod CREATE requests, mutatePods from the decoded Pod
 * on (admission-webhook/main.go:616-704): the exclude / mirror-pod checks
 * :625-633, filterPodDefaults :72-97 (label selectors with matchLabels and
 * matchExpressions In / NotIn / Exists / DoesNotExist, then the same-namespace
 * test), safeToApplyPodDefaultsOnPod :101-166 with mergeImagePullSecrets :170,
 * mergeEnv :217, mergeEnvFrom :254, mergeVolumeMounts :266, mergeVolumes :321,
 * mergeContainers :364, mergeTolerations :411, mergeMap :454,
 * applyPodDefaultsOnPod :480-556, applyPodDefaultsOnContainer :561-578,
 * setCommandAndArgs :582-597 and the patch of :685-701.
 * pods: one Pod per request.  poddefaults: every PodDefault the caller has
 * listed (the reference lists per request, :635-637; a caller keeps them
 * cached), grouped so that ranges[2i], ranges[2i+1] = the [begin, end) of the
 * PodDefaults in Pod i's namespace (NULL = all of them for every Pod).
 * Output per Pod: an RFC 6902 patch replacing /metadata and /spec (two "add"
 * operations), or length 0 = allowed without patch.  The reference's patch has
 * Go-map operation order; equality is by effect (apply it, compare the Pod).
 * status: NBRM_E_CONFLICT = rejected, conflict while merging (:671-679);
 * NBRM_E_SELECTOR = rejected, bad selector (:653-656).  result = number of
 * PodDefaults applied.  Limits: 32 PodDefaults applied to one Pod, 32 items
 * appended to one list (NBRM_E_LIMIT beyond). */
int nbrm_poddefault_mutate(nbrm_ctx* ctx, const nbrm_docs* pods, const nbrm_docs* poddefaults,
                           const uint32_t* ranges, nbrm_out* out);
int nbrm_poddefault_mutate_resident(nbrm_ctx* ctx);

/* ---- profile-controller desired objects ---------------------------------------
 * Replaces, for a batch of Profiles, the object construction inside
 * ProfileReconciler.Reconcile (profile-controller/controllers/
 * profile_controller.go:105-334): the Namespace :127-137 with setNamespaceLabels
 * :754-772 applied to the configured labels file (nbrm_env.namespace_labels_json),
 * getAuthorizationPolicy :419-504 with its metadata :512-520, the two
 * ServiceAccounts and their RoleBindings (updateServiceAccount :592-637), the owner
 * RoleBinding namespaceAdmin :230-246 and the ResourceQuota :253-262, each with the
 * controller ownerReference to the Profile.  Output: 8 objects per Profile in that
 * order; the ResourceQuota has length 0 when spec.resourceQuotaSpec.hard is empty.
 * Plugins (AWS IRSA / GCP workload identity, :286-294) and the finalizer are I/O
 * and bookkeeping and stay with the caller.  The Istio proto JSON shape of the
 * AuthorizationPolicy is recalled, not pinned by a reference test. */
int nbrm_profile_generate(nbrm_ctx* ctx, const nbrm_docs* profiles, nbrm_out* out);
int nbrm_profile_generate_resident(nbrm_ctx* ctx);

/* ---- tensorboard-controller desired objects -----------------------------------
 * Replaces, for a batch of Tensorboard CRs, generateDeployment / generateService /
 * generateVirtualService and the three ctrl.SetControllerReference calls of
 * TensorboardReconciler.Reconcile (tensorboard-controller/controllers/
 * tensorboard_controller.go:67-160, :167-378).  node_names (may be NULL): per Tensorboard the
 * node a running Pod mounts its ReadWriteOnce PVC on, as a JSON string ("" or an empty
 * document: none) — what the PVC Get and the Pod List of the RWO_PVC_SCHEDULING path
 * (:207-232, :427-468) lead to; used only when that variable is "true".  flags:
 * NBRM_GEN_OWNER_REF.  Output: Deployment, Service, VirtualService per Tensorboard.
 * status: NBRM_E_ENV when TENSORBOARD_IMAGE, ISTIO_GATEWAY or ISTIO_HOST is not set. */
int nbrm_tensorboard_generate(nbrm_ctx* ctx, const nbrm_docs* tensorboards, const nbrm_docs* node_names,
                              uint32_t flags, nbrm_out* out);

/* After the Get of the Deployment (tensorboard_controller.go:484-510, :119-152): per Tensorboard
 * the desired Deployment (output 3 i of nbrm_tensorboard_generate) and the found one (empty =
 * NotFound: the caller Creates the desired object and nothing is returned).  Output 2 i: the
 * body of the Update — the found object with the labels, replicas and affinity of the desired
 * one (CopyDeploymentSetFields :512-533) — or length 0 when nothing differs; output 2 i + 1: the
 * Tensorboard's status as r.Status().Update sends it (the Deployment's first condition appended
 * when its type is not the last deploymentState, readyReplicas).  result: 1 = Update needed. */
int nbrm_tensorboard_update(nbrm_ctx* ctx, const nbrm_docs* tensorboards, const nbrm_docs* desired,
                            const nbrm_docs* found, nbrm_out* out);

/* ---- pvcviewer-controller: what a reconcile writes ------------------------------
 * Replaces, for a batch of PVCViewer CRs, the bodies reconcileDeployment / reconcileService /
 * reconcileVirtualService build between their Get and their Create or Update
 * (pvcviewer-controller/controllers/pvcviewer_controller.go:149-336), ownerReference included
 * (SetControllerReference :198,:240,:326).  found (may be NULL: every Get was NotFound): 3 n
 * documents, 3 i + k = the Deployment / Service / VirtualService the Get returned for viewer i,
 * empty = NotFound; a found object keeps everything the reconciler does not overwrite.
 * node_names (may be NULL): per viewer the node its ReadWriteOnce PVC is mounted on, as a JSON
 * string — the outcome of generateAffinity's PVC Get and Pod List (:372-443); used for
 * rwoScheduling viewers whose Deployment is being created (:170).  Output: 3 objects per
 * viewer; Service and VirtualService have length 0 when spec.networking is the zero value
 * (:212,:253).  status: NBRM_E_CONFLICT when a found object has another controller. */
int nbrm_pvcviewer_reconcile(nbrm_ctx* ctx, const nbrm_docs* viewers, const nbrm_docs* found,
                             const nbrm_docs* node_names, nbrm_out* out);

/* The viewer's defaulting and validating webhook: (*PVCViewer).Default and validate
 * (pvcviewer-controller/api/v1alpha1/pvcviewer_webhook.go:70-146, :153-178).  Output: one object
 * per viewer, the viewer as Default leaves it.  result: 0 = valid, 1 = "PVC name must be
 * specified", 2 = "PodSpec must be specified", 3 = "PVC <pvc> must be used in the podSpec". */
int nbrm_pvcviewer_webhook(nbrm_ctx* ctx, const nbrm_docs* viewers, nbrm_out* out);

/* reconcileStatus (pvcviewer_controller.go:339-369): per viewer its Deployment (empty = the Get
 * failed: ready false).  Output: the status the reconciler sends — url from the networking,
 * ready = replicas == readyReplicas, the Deployment's first condition appended when the last
 * one differs.  NBRM_E_NIL_REPLICAS where the reference dereferences a nil Spec.Replicas. */
int nbrm_pvcviewer_status(nbrm_ctx* ctx, const nbrm_docs* viewers, const nbrm_docs* deployments, nbrm_out* out);

/* ---- composing generators --------------------------------------------------
 * Run one generator on arbitrary slots: tuple i = document i of slot_a (and of
 * slot_b / slot_c when >= 0).  Generators: 0 notebook-controller objects (3
 * per Notebook), 1 ODH objects (7), 2 create-or-update decision (slot_a desired,
 * slot_b found), 3 status (Notebook, StatefulSet, Pod), 4 culling pass
 * (Notebook, kernels, terminals; arg0/arg1 = idle / period minutes), 5 webhook
 * mutation (flags NBRM_MUT_*), 6 restart guard (mutated, request, old),
 * 7 PodDefault admission (Pods, shared PodDefault arena), 8 Profile objects (8),
 * 9 ODH write decision (slot_a desired ODH objects, slot_b found; result word 0 nothing,
 * 1 create, 2 update: reconcileRoute / reconcileNetworkPolicy / reconcileRoleBinding and the
 * create-once ServiceAccount, Service, Secret), 10 JSON diff (RFC 6902), 11 / 12 AdmissionReview open /
 * close, 13 Notebook version conversion, 14 tensorboard objects (3), 15 what a PVCViewer reconcile
 * writes (3; slot_b = the shared arena of found objects), 16 the viewer's webhook, 17 tensorboard
 * Deployment update + status (2), 18 the viewer's status.  use_side_input: pass
 * the bytes of nbrm_upload_bytes (seeds / pod_exists / ranges).  The result is
 * read with nbrm_download or moved on with nbrm_result_to_slot. */
int nbrm_run(nbrm_ctx* ctx, int generator, int slot_a, int slot_b, int slot_c, uint32_t flags,
             uint64_t now_unix, uint32_t arg0, uint32_t arg1, int use_side_input);
/* Copy the documents of a slot to host memory (status and result are NULL). */
int nbrm_download_slot(nbrm_ctx* ctx, int slot, nbrm_out* out);

/* ---- uploads that overlap the running generators -----------------------------
 * A controller drains its workqueue while the previous batch is still being
 * reconciled (controller-runtime's workers run concurrently, MaxConcurrentReconciles);
 * the batch equivalent is to copy the next batch of requests to the device while
 * the current one runs.  nbrm_upload_async is nbrm_upload on a stream of its own:
 * generators that read `slot` afterwards wait for the copy on the device, and the
 * copy waits for the generators already launched that read the slot's previous
 * documents — so alternate two slots between consecutive batches.  docs->data and
 * docs->off must be pinned (nbrm_host_alloc) and stay untouched until a generator
 * that reads the slot has been run and nbrm_sync (or a download) has returned. */
int nbrm_upload_async(nbrm_ctx* ctx, int slot, const nbrm_docs* docs);

/* ---- downloads that overlap the next generator ------------------------------
 * The reference hands each object to the API server as soon as it is built
 * (r.Create / r.Update, notebook_controller.go:160,176,198,208 and the ODH
 * reconcilers) and goes on with the next one; the batch equivalent is to send a
 * finished result to the host while the device already runs the next generator.
 * nbrm_download_begin takes the current result away from the context (like
 * nbrm_result_to_slot it is then no longer "the result") and starts copying it
 * on a second stream into pinned buffers that belong to download lane `lane`
 * (0 .. NBRM_DL_LANES-1); nbrm_download_slot_begin does the same for the
 * documents of an input slot, which stay where they are.  nbrm_download_wait
 * blocks until that lane's copy is complete and fills *out exactly as
 * nbrm_download / nbrm_download_slot would (same return value: 1 when a status
 * word is non-zero).  The pointers stay valid until the next *_begin on the
 * same lane.  A lane holds one download at a time: begin on a lane whose last
 * download was not waited for is an error (-2). */
#define NBRM_DL_LANES 4
int nbrm_download_begin(nbrm_ctx* ctx, int lane);
int nbrm_download_slot_begin(nbrm_ctx* ctx, int slot, int lane);
int nbrm_download_wait(nbrm_ctx* ctx, int lane, nbrm_out* out);

/* ---- Notebook version conversion ------------------------------------------------
 * ConvertTo / ConvertFrom of the v1 and v1alpha1 spokes around the v1beta1 hub
 * (notebook-controller/api/v1/notebook_conversion.go:25-69, api/v1alpha1 likewise).
 * to_version: 0 kubeflow.org/v1beta1, 1 kubeflow.org/v1, 2 kubeflow.org/v1alpha1.
 * As in the reference only spec.template.spec, status.readyReplicas,
 * status.containerState and each condition's type / lastProbeTime / reason /
 * message cross over: a condition's status comes out "" and its
 * lastTransitionTime is dropped (:30-38, :56-64).  The reference's functions do
 * not touch ObjectMeta; NBRM_CONVERT_KEEP_METADATA carries the source's through. */
enum { NBRM_CONVERT_KEEP_METADATA = 1 };
int nbrm_notebook_convert(nbrm_ctx* ctx, const nbrm_docs* notebooks, int to_version, uint32_t flags,
                          nbrm_out* out);

/* ---- wire formats of the two webhooks -----------------------------------------
 * nbrm_json_diff: the RFC 6902 patch that turns originals[i] into mutated[i]
 * (length 0 when they are equal) — what admission.PatchResponseFromRaw
 * (odh-notebook-controller/controllers/notebook_webhook.go:301) and
 * jsonpatch.CreatePatch (admission-webhook/main.go:693) compute.  Both Go
 * libraries order their operations by map iteration, so a patch is defined by
 * its effect: applied to the original it gives the mutated document.  Arrays
 * that only grew get one "add .../-" per new element.
 *
 * nbrm_notebook_webhook_review: NotebookWebhook.Handle end to end for a batch of
 * AdmissionReview requests (notebook_webhook.go:234-302 and controller-runtime's
 * response writer): request.object is mutated as by nbrm_odh_mutate (flags
 * NBRM_MUT_*; is_update as there, with request.oldObject), and each output is
 * the AdmissionReview response: kind / apiVersion echoed, response.uid echoed,
 * allowed, patch (base64 of the JSON patch) and patchType when there is one;
 * admission.Denied as allowed:false with status code 403 and the reference's
 * message.  The caller sends reviews of one operation per call.
 *
 * nbrm_poddefault_review: serve(mutatePods) (admission-webhook/main.go:599-749)
 * for a batch of AdmissionReview requests: the Pod's namespace falls back to
 * request.namespace (:617-620), PodDefaults are applied as by
 * nbrm_poddefault_mutate, and each output is {"response":{"uid":..,"allowed":..
 * [,"patch":<base64>,"patchType":"JSONPatch"]}}; a conflict or a bad selector
 * gives allowed:false with a status message (the reference's text carries
 * Go-syntax dumps and is not reproduced); a request for anything but core/v1
 * pods gives "{}" — a review without a response, as the reference writes it
 * (:602-605, :733-737).  The Content-Type check of serve (:718-721: no body at
 * all on a mismatch) is the HTTP server's and stays with the caller. */
int nbrm_json_diff(nbrm_ctx* ctx, const nbrm_docs* originals, const nbrm_docs* mutated, nbrm_out* out);
int nbrm_notebook_webhook_review(nbrm_ctx* ctx, const nbrm_docs* reviews, uint32_t flags, int is_update,
                                 nbrm_out* out);
int nbrm_poddefault_review(nbrm_ctx* ctx, const nbrm_docs* reviews, const nbrm_docs* poddefaults,
                           const uint32_t* ranges, nbrm_out* out);

/* ---- resident documents: index once ------------------------------------------
 * No counterpart in the reference as a function; it is what the reference's
 * informer cache does for it: objects that stay around (the found StatefulSets,
 * Services, Pods ... a reconcile compares against) are decoded when they arrive,
 * not on every reconcile.  nbrm_slot_index runs the structural index pass
 * (kubeflow-rm_b200/csrc/nbrm_tape.cuh) over the documents of a slot and keeps
 * the result in device memory beside them; every later generator that reads
 * the slot hops through that index instead of scanning the text.  Results are
 * byte-identical with and without it.  The index is dropped when the slot's
 * contents change (nbrm_upload, nbrm_upload_async, nbrm_result_to_slot);
 * calling this again on an indexed slot does nothing.  Cost: about one pass
 * over the slot's bytes, plus device memory of 1.6x the slot's bytes. */
int nbrm_slot_index(nbrm_ctx* ctx, int slot);

/* ---- inspection: the structural index --------------------------------------
 * No counterpart in the reference.  The plan kernels index every input document
 * before they walk it (kubeflow-rm_b200/csrc/nbrm_tape.cuh: position, kind and
 * nesting depth of each { } [ ] : , outside strings, plus the links the walkers
 * hop through).  This call runs that index pass alone over the documents of a
 * slot and copies the result to host memory, so it can be checked against a
 * scalar pass over the same text.
 *   entries  n * cap 64-bit words (low half: pos | kind << 21 | depth << 25,
 *            high half: the two 16-bit links), document i at entries[i * cap]
 *   counts   n entry counts (0: the document has no index — over 2 MiB, nested
 *            deeper than 63, or more than cap entries — and is walked as text)
 *   flags    n words, bit 0: whitespace occurs between tokens */
int nbrm_tape_dump(nbrm_ctx* ctx, int slot, uint32_t cap, uint64_t* entries, uint32_t* counts,
                   uint32_t* flags);

/* ---- offsets without allocation -----------------------------------------
 * No counterpart in the reference: this entry point exists for embedders, not
 * to replace a reference function.
 *
 * Exclusive prefix sum of n object lengths: d_off[i] = d_lens[0] + ... +
 * d_lens[i-1], d_off[n] = total.  It is the offset step every generator above
 * runs between planning and writing its objects, offered on its own for
 * callers that lay out packed batches themselves but must not allocate device
 * memory while doing so: an admission or reconcile worker that was given a
 * fixed slice of a GPU shared with the users' notebook pods, where a
 * cudaMalloc in the request path can fail under the neighbours' memory
 * pressure and, even when it succeeds, stalls every stream on the device.
 * This call uses no cudaMalloc/cudaFree and no allocating library routine; its
 * only scratch is a 32 KiB array (4096 tile sums) that lives inside the loaded
 * library on every GPU that loads it, until the process exits.
 *
 *   d_lens, d_off  DEVICE pointers owned by the caller (n and n+1 entries).
 *   stream         a cudaStream_t (NULL = the default stream); the call only
 *                  enqueues work, it does not wait for it.
 *   err, err_cap   optional buffer for a message when the result is not 0.
 *
 * Limit: n <= NBRM_SCAN_NOALLOC_MAX_N (4 194 304); larger n is rejected with
 * -3 and nothing is launched.  The device used is the calling thread's
 * current device.  The first call on each device also looks up the scratch
 * array's device address (cudaGetSymbolAddress); make that first call before
 * starting any stream capture.  All calls on a device share the scratch, so
 * two calls that can overlap on the GPU (different streams, no ordering
 * between them) race: order them with an event or use one stream. */
#define NBRM_SCAN_NOALLOC_MAX_N 4194304u
int nbrm_scan_lengths_noalloc(const uint32_t* d_lens, uint64_t n, uint64_t* d_off, void* stream,
                              char* err, size_t err_cap);

#ifdef __cplusplus
}
#endif
#endif /* NBRM_H_ */
