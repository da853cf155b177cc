// notebook_reconciler.cc — see notebook_reconciler.h.
#include "notebook_reconciler.h"

#include <ctype.h>

#include "../csrc/nbrm_json.cuh"

namespace nbrm_host {

namespace {
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

const char* status_text(uint32_t st) {
  switch (st) {
    case NBRM_E_PARSE: return "object is not valid JSON";
    case NBRM_E_NO_CONTAINERS: return "Notebook has no containers (the reference indexes Containers[0])";
    case NBRM_E_PORTS_EMPTY: return "containers[0].ports is empty (the reference indexes Ports[0])";
    case NBRM_E_NIL_REPLICAS: return "found StatefulSet has no spec.replicas";
    case NBRM_E_SEGS: return "object too fragmented for the plan";
    default: return "rejected";
  }
}

std::string current_status(const std::string& nb) {
  nbrm::Doc d;
  d.p = nb.data();
  d.n = (uint32_t)nb.size();
  nbrm::Span s = nbrm::find(d, nbrm::root(d), "status");
  return std::string(d.p + s.off, s.len);
}
}  // namespace

int NotebookReconciler::ReconcileBatch(const std::vector<Request>& reqs, uint64_t now_unix,
                                       BatchStats* stats, std::vector<std::string>* errors) {
  BatchStats st;
  st.keys = reqs.size();
  std::vector<std::string> errs(reqs.size());
  // r.Get(Notebook) :126; gone -> nil :128; terminating -> nil :135-137
  std::vector<size_t> live;
  std::vector<std::string> nbs;
  for (size_t i = 0; i < reqs.size(); ++i) {
    std::string ev;
    if (store_->Get(Key{"Event", reqs[i].ns, reqs[i].name}, &ev) == Err::Ok) {  // :95-119
      // nbNameFromInvolvedObject :666-690
      const std::string kind = read_string(ev, {"involvedObject", "kind"});
      const std::string oname = read_string(ev, {"involvedObject", "name"});
      const std::string ons = read_string(ev, {"involvedObject", "namespace"});
      std::string nb_name, pod;
      if (kind == "StatefulSet") {
        nb_name = oname;
      } else if (kind == "Pod" && store_->Get(Key{"Pod", ons, oname}, &pod) == Err::Ok) {
        nb_name = read_string(pod, {"metadata", "labels", "notebook-name"});
      }
      if (nb_name.empty()) {
        errs[i] = kind == "Pod" && pod.empty() ? "pod of the event not found" : "object isn't related to a Notebook";
        continue;
      }
      std::string nb;
      if (store_->Get(Key{"Notebook", reqs[i].ns, nb_name}, &nb) != Err::Ok) continue;  // ignoreNotFound :111
      std::string lower = kind;
      for (char& ch : lower) ch = (char)tolower((unsigned char)ch);
      // EventRecorder.Eventf(notebook, type, reason, "Reissued from %s/%s: %s") :116-117
      const std::string obj =
          "{\"apiVersion\":\"v1\",\"kind\":\"Event\",\"metadata\":{\"name\":\"" + nb_name + "." + reqs[i].name +
          "\",\"namespace\":\"" + reqs[i].ns + "\"},\"involvedObject\":{\"kind\":\"Notebook\",\"name\":\"" + nb_name +
          "\",\"namespace\":\"" + reqs[i].ns + "\"},\"type\":\"" + read_string(ev, {"type"}) + "\",\"reason\":\"" +
          read_string(ev, {"reason"}) + "\",\"message\":\"Reissued from " + lower + "/" + oname + ": " +
          read_string(ev, {"message"}) + "\"}";
      Err e = store_->Create(obj);
      if (e == Err::Ok || e == Err::AlreadyExists) ++st.events_reissued;
      continue;
    }
    std::string nb;
    if (store_->Get(Key{"Notebook", reqs[i].ns, reqs[i].name}, &nb) != Err::Ok) {
      ++st.not_found;
      continue;
    }
    if (read_meta(nb).has_deletion_timestamp) {
      ++st.terminating;
      continue;
    }
    live.push_back(i);
    nbs.push_back(std::move(nb));
  }
  if (!live.empty()) {
    Packed p_nb, p_found, p_sts, p_pod;
    std::vector<bool> pod_found(live.size());
    for (size_t j = 0; j < live.size(); ++j) {
      const Request& r = reqs[live[j]];
      p_nb.add(nbs[j]);
      std::string sts, svc, vs, pod;
      if (store_->Get(Key{"StatefulSet", r.ns, r.name}, &sts) != Err::Ok) sts.clear();      // :147
      if (store_->Get(Key{"Service", r.ns, r.name}, &svc) != Err::Ok) svc.clear();          // :189
      if (use_istio_ &&
          store_->Get(Key{"VirtualService", r.ns, "notebook-" + r.ns + "-" + r.name}, &vs) != Err::Ok)
        vs.clear();                                                                          // :636
      pod_found[j] = store_->Get(Key{"Pod", r.ns, r.name + "-0"}, &pod) == Err::Ok;         // :221
      if (!pod_found[j]) pod.clear();
      p_found.add(sts);
      p_found.add(svc);
      p_found.add(vs);
      p_sts.add(sts);  // foundStateful as read before any write (:145-147, used at :229)
      p_pod.add(pod);
    }
    const uint32_t flags = NBRM_GEN_STS | NBRM_GEN_SVC | (use_istio_ ? NBRM_GEN_VS : 0) | NBRM_GEN_OWNER_REF;
    nbrm_docs d_nb = p_nb.docs(), d_found = p_found.docs(), d_sts = p_sts.docs(), d_pod = p_pod.docs();
    nbrm_out out;
    int rc = nbrm_nbc_reconcile(ctx_, &d_nb, &d_found, flags, &out);
    if (rc < 0) {
      for (size_t j = 0; j < live.size(); ++j) errs[live[j]] = nbrm_last_error(ctx_);
      st.errors += live.size();
    } else {
      for (size_t j = 0; j < live.size(); ++j) {
        for (int k = 0; k < 3; ++k) {
          const size_t o = 3 * j + k;
          if (out.status[o]) {
            errs[live[j]] = status_text(out.status[o]);
            continue;
          }
          const uint64_t a = out.off[o], b = out.off[o + 1];
          if (a == b) continue;  // nothing to write
          std::string obj(out.data + a, b - a);
          const bool found = d_found.off[o + 1] > d_found.off[o];
          Err e = found ? store_->Update(obj) : store_->Create(obj);  // :174,:205,:653 / :151,:192,:641
          if (e != Err::Ok) errs[live[j]] = found ? "update failed" : "create failed";
          else if (found) ++st.updates;
          else ++st.creates;
        }
      }
      // updateNotebookStatus :274-288
      rc = nbrm_notebook_status(ctx_, &d_nb, &d_sts, &d_pod, now_unix, &out);
      if (rc < 0) {
        for (size_t j = 0; j < live.size(); ++j) errs[live[j]] = nbrm_last_error(ctx_);
      } else {
        for (size_t j = 0; j < live.size(); ++j) {
          if (!errs[live[j]].empty()) continue;  // the reference returned before reaching :229
          const Request& r = reqs[live[j]];
          std::string status(out.data + out.off[j], out.off[j + 1] - out.off[j]);
          if (write_if_changed_ && status == current_status(nbs[j])) {
            ++st.status_unchanged;
          } else if (store_->UpdateStatus(Key{"Notebook", r.ns, r.name}, status) == Err::Ok) {
            ++st.status_updates;
          } else {
            errs[live[j]] = "status update failed";
          }
        }
      }
      // restart annotation :237-269
      for (size_t j = 0; j < live.size(); ++j) {
        if (!errs[live[j]].empty()) continue;
        if (read_annotation(nbs[j], "notebooks.opendatahub.io/notebook-restart") != "\"true\"") continue;
        const Request& r = reqs[live[j]];
        if (store_->Delete(Key{"Pod", r.ns, r.name + "-0"}) != Err::Ok) {
          errs[live[j]] = "pod delete failed";  // :254-256: Delete of a pod that was not found errors
          continue;
        }
        std::string cur;
        if (store_->Get(Key{"Notebook", r.ns, r.name}, &cur) == Err::Ok &&
            store_->Update(without_annotation(cur, "notebooks.opendatahub.io/notebook-restart")) == Err::Ok)
          ++st.pod_restarts;
        else
          errs[live[j]] = "notebook update failed";
      }
    }
  }
  for (size_t i = 0; i < reqs.size(); ++i)
    if (!errs[i].empty()) ++st.errors;
  st.reconciled = live.size();
  if (stats) *stats = st;
  if (errors) *errors = std::move(errs);
  return 0;
}

Result NotebookReconciler::Reconcile(const Request& req, uint64_t now_unix, std::string* error) {
  std::vector<std::string> errs;
  ReconcileBatch({req}, now_unix, nullptr, &errs);
  if (error) *error = errs.empty() ? std::string() : errs[0];
  return Result();  // the reference never asks for a timed requeue here (:271)
}

}  // namespace nbrm_host
