// odh_reconciler.cc — see odh_reconciler.h.
#include "odh_reconciler.h"

#include <random>

#include "../csrc/nbrm_json.cuh"

namespace nbrm_host {

namespace {
const int GEN_ODH_DECIDE = 9;  // nbrm_run generator id (include/nbrm.h)
const int PER_NB = 7;          // ctrl-np, oauth-np, route, sa, service, secret, rolebinding

struct Packed {
  std::string data;
  std::vector<uint64_t> off{0};
  void add(const std::string& s) {
    data += s;
    off.push_back(data.size());
  }
  nbrm_docs docs() const {
    nbrm_docs d;
    d.data = data.data();
    d.off = off.data();
    d.n = off.size() - 1;
    return d;
  }
};

// Kind and name of ODH object k of Notebook `name` (notebook_network.go:142,183,
// notebook_route.go:41, notebook_oauth.go:52,119,200, notebook_rbac.go:147).
Key object_key(int k, const std::string& ns, const std::string& name) {
  switch (k) {
    case 0: return Key{"NetworkPolicy", ns, name + "-ctrl-np"};
    case 1: return Key{"NetworkPolicy", ns, name + "-oauth-np"};
    case 2: return Key{"Route", ns, name};
    case 3: return Key{"ServiceAccount", ns, name};
    case 4: return Key{"Service", ns, name + "-tls"};
    case 5: return Key{"Secret", ns, name + "-oauth-config"};
    default: return Key{"RoleBinding", ns, "elyra-pipelines-" + name};
  }
}

bool has_pull_secret(const std::string& sa) {
  nbrm::Doc d;
  d.p = sa.data();
  d.n = (uint32_t)sa.size();
  const nbrm::Span ips = nbrm::find(d, nbrm::root(d), "imagePullSecrets");
  return nbrm::is_array(d, ips) && !nbrm::is_empty_container(d, ips);
}
}  // namespace

int OpenshiftNotebookReconciler::ReconcileBatch(const std::vector<Request>& reqs, OdhStats* stats,
                                                std::vector<std::string>* errors) {
  OdhStats st;
  st.keys = reqs.size();
  std::vector<std::string> errs(reqs.size());
  std::vector<size_t> live;
  std::vector<std::string> nbs;
  for (size_t i = 0; i < reqs.size(); ++i) {
    std::string nb;
    if (store_->Get(Key{"Notebook", reqs[i].ns, reqs[i].name}, &nb) != Err::Ok) {  // :158-161
      ++st.not_found;
      continue;
    }
    live.push_back(i);
    nbs.push_back(std::move(nb));
  }
  const uint32_t flags = NBRM_GEN_OWNER_REF | NBRM_GEN_OWNER_V1 | (rbac_ ? NBRM_ODH_RBAC : 0);
  if (!live.empty()) {
    Packed p_nb;
    for (auto& nb : nbs) p_nb.add(nb);
    // cookie seeds: the reference draws crypto/rand bytes per Secret (notebook_oauth.go:192-195)
    std::vector<uint8_t> seeds(16 * live.size());
    std::random_device rd;
    for (auto& b : seeds) b = (uint8_t)rd();
    nbrm_docs d_nb = p_nb.docs();
    nbrm_out out;
    int rc = nbrm_odh_generate(ctx_, &d_nb, seeds.data(), flags, &out);
    if (rc < 0) {
      for (size_t j = 0; j < live.size(); ++j) errs[live[j]] = nbrm_last_error(ctx_);
    } else {
      // keep the desired objects: the context's host buffers are reused by the next call
      Packed desired;
      std::vector<uint32_t> gen_status(out.status, out.status + live.size());
      for (size_t o = 0; o < live.size() * PER_NB; ++o)
        desired.add(std::string(out.data + out.off[o], out.off[o + 1] - out.off[o]));
      // what each reconciler Gets
      auto decide = [&](const Packed& want, const Packed& found, nbrm_out* res) {
        nbrm_docs a = want.docs(), b = found.docs();
        if (nbrm_upload(ctx_, 0, &a) || nbrm_upload(ctx_, 1, &b)) return -1;
        if (nbrm_run(ctx_, GEN_ODH_DECIDE, 0, 1, -1, 0, 0, 0, 0, 0)) return -1;
        return nbrm_download(ctx_, res) < 0 ? -1 : 0;
      };
      Packed found;
      for (size_t j = 0; j < live.size(); ++j) {
        const Request& r = reqs[live[j]];
        for (int k = 0; k < PER_NB; ++k) {
          std::string f;
          if (store_->Get(object_key(k, r.ns, r.name), &f) != Err::Ok) f.clear();
          found.add(f);
        }
      }
      nbrm_out dec;
      if (decide(desired, found, &dec) != 0) {
        for (size_t j = 0; j < live.size(); ++j) errs[live[j]] = nbrm_last_error(ctx_);
      } else {
        // copy the decisions out before any retry reuses the buffers
        std::vector<std::string> writes(live.size() * PER_NB);
        std::vector<uint32_t> what(dec.result, dec.result + live.size() * PER_NB);
        std::vector<uint32_t> dstat(dec.status, dec.status + live.size() * PER_NB);
        for (size_t o = 0; o < writes.size(); ++o) writes[o].assign(dec.data + dec.off[o], dec.off[o + 1] - dec.off[o]);
        for (size_t j = 0; j < live.size(); ++j) {
          const Request& r = reqs[live[j]];
          if (gen_status[j]) {
            errs[live[j]] = "the Notebook cannot be handled";
            continue;
          }
          for (int k = 0; k < PER_NB && errs[live[j]].empty(); ++k) {
            const size_t o = j * PER_NB + k;
            if (dstat[o]) {
              errs[live[j]] = "compare failed";
              break;
            }
            if (what[o] == 1) {  // NotFound -> Create
              Err e = store_->Create(writes[o]);
              if (e == Err::Ok) ++st.creates;
              else if (e != Err::AlreadyExists) errs[live[j]] = "create failed";
            } else if (what[o] == 2) {  // retry.RetryOnConflict(retry.DefaultRetry, ...)
              std::string obj = writes[o];
              Err e = store_->Update(obj);
              for (int attempt = 1; e == Err::Conflict && attempt < 5; ++attempt) {
                ++st.conflicts_retried;
                std::string f;
                if (store_->Get(object_key(k, r.ns, r.name), &f) != Err::Ok) break;
                Packed w1, f1;
                w1.add(desired.data.substr(desired.off[o], desired.off[o + 1] - desired.off[o]));
                f1.add(f);
                nbrm_out one;
                if (decide(w1, f1, &one) != 0) break;
                if (one.result[0] != 2) {  // someone else brought it in line meanwhile
                  e = Err::Ok;
                  obj.clear();
                  break;
                }
                obj.assign(one.data + one.off[0], one.off[1] - one.off[0]);
                e = store_->Update(obj);
              }
              if (e == Err::Ok) {
                if (!obj.empty()) ++st.updates;
              } else {
                errs[live[j]] = "update failed";
              }
            }
          }
        }
      }
    }
    // RemoveReconciliationLock :118-146, when ReconciliationLockIsEnabled :108-115
    for (size_t j = 0; j < live.size(); ++j) {
      if (!errs[live[j]].empty()) continue;
      const Request& r = reqs[live[j]];
      if (read_annotation(nbs[j], "kubeflow-resource-stopped") != "\"odh-notebook-controller-lock\"") continue;
      const double waits[2] = {1.0, 5.0};
      for (int attempt = 0; attempt < 3; ++attempt) {
        std::string sa;
        if (store_->Get(Key{"ServiceAccount", r.ns, r.name}, &sa) == Err::Ok && has_pull_secret(sa)) break;
        if (attempt < 2) {
          ++st.pull_secret_waits;
          if (sleep_) sleep_(waits[attempt]);
        }
      }
      if (store_->Patch(Key{"Notebook", r.ns, r.name},
                        "{\"metadata\":{\"annotations\":{\"kubeflow-resource-stopped\":null}}}") == Err::Ok)
        ++st.locks_removed;
      else
        errs[live[j]] = "lock patch failed";
    }
  }
  for (auto& e : errs)
    if (!e.empty()) ++st.errors;
  st.reconciled = live.size();
  if (stats) *stats = st;
  if (errors) *errors = std::move(errs);
  return 0;
}

}  // namespace nbrm_host
