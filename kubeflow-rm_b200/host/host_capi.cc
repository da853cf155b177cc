// host_capi.cc — extern "C" wrappers declared in include/nbrm_host.h.
#include <stdlib.h>
#include <string.h>

#include "../../include/nbrm_host.h"
#include "notebook_reconciler.h"
#include "odh_reconciler.h"
#include "profile_reconciler.h"
#include "workqueue.h"

using namespace nbrm_host;

struct nbrm_store {
  Store s;
};
struct nbrm_reconciler {
  NotebookReconciler r;
};
struct nbrm_odh_reconciler {
  double slept = 0;  // seconds the lock wait asked for (simulated: nothing sleeps)
  OpenshiftNotebookReconciler r;
  nbrm_odh_reconciler(Store* s, nbrm_ctx* c, bool rbac)
      : r(s, c, rbac, [this](double sec) { slept += sec; }) {}
};
struct nbrm_workqueue {
  WorkQueue q;
};

static std::vector<Request> parse_keys(const char* keys) {
  std::vector<Request> reqs;
  const char* p = keys;
  while (p && *p) {
    const char* e = strchr(p, '\n');
    std::string line = e ? std::string(p, e - p) : std::string(p);
    size_t slash = line.find('/');
    if (slash != std::string::npos) reqs.push_back(Request{line.substr(0, slash), line.substr(slash + 1)});
    if (!e) break;
    p = e + 1;
  }
  return reqs;
}

extern "C" {

nbrm_store* nbrm_store_new(void) { return new nbrm_store(); }
void nbrm_store_free(nbrm_store* s) { delete s; }
int nbrm_store_create(nbrm_store* s, const char* json, uint64_t len) {
  return (int)s->s.Create(std::string(json, len));
}
int nbrm_store_update(nbrm_store* s, const char* json, uint64_t len) {
  return (int)s->s.Update(std::string(json, len));
}
int nbrm_store_delete(nbrm_store* s, const char* kind, const char* ns, const char* name) {
  return (int)s->s.Delete(Key{kind, ns, name});
}
int nbrm_store_get(nbrm_store* s, const char* kind, const char* ns, const char* name, char** out,
                   uint64_t* len) {
  std::string o;
  Err e = s->s.Get(Key{kind, ns, name}, &o);
  if (e != Err::Ok) return (int)e;
  *out = (char*)malloc(o.size() + 1);
  memcpy(*out, o.data(), o.size());
  (*out)[o.size()] = 0;
  *len = o.size();
  return 0;
}
void nbrm_store_counters(nbrm_store* s, uint64_t* c, int reset) {
  Counters k = s->s.counters();
  c[0] = k.gets; c[1] = k.creates; c[2] = k.updates; c[3] = k.status_updates; c[4] = k.deletes; c[5] = k.conflicts;
  if (reset) s->s.reset_counters();
}
uint64_t nbrm_store_size(nbrm_store* s) { return s->s.size(); }
void nbrm_store_free_string(char* p) { free(p); }

int nbrm_store_patch(nbrm_store* s, const char* kind, const char* ns, const char* name, const char* patch,
                     uint64_t len, int json_patch) {
  const Key k{kind, ns, name};
  return (int)(json_patch ? s->s.PatchJson(k, std::string(patch, len)) : s->s.Patch(k, std::string(patch, len)));
}

nbrm_odh_reconciler* nbrm_odh_reconciler_new(nbrm_store* s, nbrm_ctx* ctx, int rbac) {
  return new nbrm_odh_reconciler(&s->s, ctx, rbac != 0);
}
void nbrm_odh_reconciler_free(nbrm_odh_reconciler* r) { delete r; }
int nbrm_odh_reconciler_reconcile(nbrm_odh_reconciler* r, const char* keys, uint64_t* stats, double* slept_seconds) {
  OdhStats st;
  std::vector<std::string> errs;
  r->slept = 0;
  r->r.ReconcileBatch(parse_keys(keys), &st, &errs);
  if (stats) {
    stats[0] = st.keys; stats[1] = st.reconciled; stats[2] = st.not_found; stats[3] = st.creates;
    stats[4] = st.updates; stats[5] = st.conflicts_retried; stats[6] = st.locks_removed;
    stats[7] = st.pull_secret_waits; stats[8] = st.errors;
  }
  if (slept_seconds) *slept_seconds = r->slept;
  return (int)st.errors;
}

struct nbrm_profile_reconciler {
  ProfileReconciler r;
};
nbrm_profile_reconciler* nbrm_profile_reconciler_new(nbrm_store* s, nbrm_ctx* ctx, const char* workload_identity,
                                                     const char* default_labels_json) {
  return new nbrm_profile_reconciler{ProfileReconciler(&s->s, ctx, workload_identity ? workload_identity : "",
                                                       default_labels_json ? default_labels_json : "")};
}
void nbrm_profile_reconciler_free(nbrm_profile_reconciler* r) { delete r; }
int nbrm_profile_reconciler_reconcile(nbrm_profile_reconciler* r, const char* names, uint64_t* stats) {
  std::vector<std::string> keys;
  const char* p = names;
  while (p && *p) {
    const char* e = strchr(p, '\n');
    std::string line = e ? std::string(p, e - p) : std::string(p);
    if (!line.empty()) keys.push_back(line);
    if (!e) break;
    p = e + 1;
  }
  ProfileStats st;
  std::vector<std::string> errs;
  r->r.ReconcileBatch(keys, &st, &errs);
  if (stats) {
    stats[0] = st.keys; stats[1] = st.reconciled; stats[2] = st.not_found; stats[3] = st.creates;
    stats[4] = st.updates; stats[5] = st.deletes; stats[6] = st.rejected; stats[7] = st.profile_writes;
    stats[8] = st.finalizers_added; stats[9] = st.finalizers_removed; stats[10] = st.plugins_applied;
    stats[11] = st.plugins_revoked; stats[12] = st.errors;
  }
  return (int)st.errors;
}

nbrm_workqueue* nbrm_workqueue_new(void) { return new nbrm_workqueue(); }
void nbrm_workqueue_free(nbrm_workqueue* q) { delete q; }
void nbrm_workqueue_add(nbrm_workqueue* q, const char* ns, const char* name) { q->q.Add(Request{ns, name}); }
void nbrm_workqueue_add_after(nbrm_workqueue* q, const char* ns, const char* name, double now, double seconds) {
  q->q.AddAfter(Request{ns, name}, now, seconds);
}
double nbrm_workqueue_add_rate_limited(nbrm_workqueue* q, const char* ns, const char* name, double now) {
  return q->q.AddRateLimited(Request{ns, name}, now);
}
void nbrm_workqueue_forget(nbrm_workqueue* q, const char* ns, const char* name) { q->q.Forget(Request{ns, name}); }
void nbrm_workqueue_done(nbrm_workqueue* q, const char* ns, const char* name) { q->q.Done(Request{ns, name}); }
uint64_t nbrm_workqueue_len(nbrm_workqueue* q) { return q->q.Len(); }
double nbrm_workqueue_next_ready(nbrm_workqueue* q) { return q->q.NextReady(); }
// keys out as "ns/name\n..." in a malloc'ed string (nbrm_store_free_string); returns the count
uint64_t nbrm_workqueue_take(nbrm_workqueue* q, uint64_t max, double now, char** out) {
  std::vector<Request> reqs;
  q->q.Take(max, now, &reqs);
  std::string text;
  for (auto& r : reqs) text += r.ns + "/" + r.name + "\n";
  *out = (char*)malloc(text.size() + 1);
  memcpy(*out, text.c_str(), text.size() + 1);
  return reqs.size();
}

nbrm_reconciler* nbrm_reconciler_new(nbrm_store* s, nbrm_ctx* ctx, int use_istio, int wic) {
  return new nbrm_reconciler{NotebookReconciler(&s->s, ctx, use_istio != 0, wic != 0)};
}
void nbrm_reconciler_free(nbrm_reconciler* r) { delete r; }
int nbrm_reconciler_reconcile(nbrm_reconciler* r, const char* keys, uint64_t now_unix, uint64_t* stats) {
  std::vector<Request> reqs;
  const char* p = keys;
  while (*p) {
    const char* e = strchr(p, '\n');
    std::string line = e ? std::string(p, e - p) : std::string(p);
    size_t slash = line.find('/');
    if (slash != std::string::npos) reqs.push_back(Request{line.substr(0, slash), line.substr(slash + 1)});
    if (!e) break;
    p = e + 1;
  }
  BatchStats st;
  std::vector<std::string> errs;
  r->r.ReconcileBatch(reqs, now_unix, &st, &errs);
  if (stats) {
    stats[0] = st.keys; stats[1] = st.reconciled; stats[2] = st.not_found; stats[3] = st.terminating;
    stats[4] = st.creates; stats[5] = st.updates; stats[6] = st.status_updates; stats[7] = st.status_unchanged;
    stats[8] = st.pod_restarts; stats[9] = st.errors; stats[10] = st.events_reissued;
  }
  return (int)st.errors;
}
}
