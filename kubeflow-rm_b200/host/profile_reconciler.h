// profile_reconciler.h — ProfileReconciler: the profile-controller's reconciler in C++ over the
// store and the GPU batch engine.
//
// Mirrors components/profile-controller/controllers/profile_controller.go: Reconcile :105-334
// (Namespace create / ownership check / label update :127-204, updateIstioAuthorizationPolicy
// :509-558, updateServiceAccount :592-637, updateRoleBinding :639-674, updateResourceQuota
// :560-590 and the delete of a quota the Profile no longer asks for :262-277,
// PatchDefaultPluginSpec :716-740, GetPluginSpec :676-714 with ApplyPlugin / RevokePlugin, the
// finalizer :292-330), appendErrorConditionAndReturn :337-347, setNamespaceLabels :754-772.
// The eight desired objects of every Profile of a batch come from one GPU pass
// (nbrm_profile_generate); what the reference then does object by object — Get, compare the
// fields that function owns, Create / Update / Delete — runs on the host against the store.
//   * A Namespace that exists and whose `owner` annotation is not the Profile's owner is not
//     taken over: the Profile gets a `Failed` condition and the reconcile ends (:198-203).
//   * PatchDefaultPluginSpec writes the Profile on EVERY reconcile, changed or not (:735): so
//     does this.
//   * Plugins (GCP workload identity, AWS IAM for service accounts) talk to cloud APIs: they are
//     callbacks, given the plugin's kind and spec; the default ones do nothing.
//   * The wait for a freshly created Namespace (backoff.Retry, 5 x 3 s :160-165) is a Get that
//     succeeds at once on this store.
#pragma once
#include <functional>
#include <string>
#include <vector>

#include "../../include/nbrm.h"
#include "store.h"

namespace nbrm_host {

struct ProfileStats {
  uint64_t keys = 0, reconciled = 0, not_found = 0, creates = 0, updates = 0, deletes = 0;
  uint64_t rejected = 0, profile_writes = 0, finalizers_added = 0, finalizers_removed = 0;
  uint64_t plugins_applied = 0, plugins_revoked = 0, errors = 0;
};

// kind, spec (JSON), the Profile (JSON); false = the plugin failed
using PluginFn = std::function<bool(const std::string&, const std::string&, const std::string&)>;

class ProfileReconciler {
 public:
  // workload_identity: the --workload-identity flag ("" = off).  default_labels_json: the
  // namespace-labels file as a JSON object of strings ("" = none); the same text must be in the
  // context's environment (nbrm_env.namespace_labels_json) for the generated Namespace.
  ProfileReconciler(Store* store, nbrm_ctx* ctx, std::string workload_identity, std::string default_labels_json,
                    PluginFn apply = nullptr, PluginFn revoke = nullptr)
      : store_(store), ctx_(ctx), wi_(std::move(workload_identity)), labels_(std::move(default_labels_json)),
        apply_(std::move(apply)), revoke_(std::move(revoke)) {}
  int ReconcileBatch(const std::vector<std::string>& names, ProfileStats* stats, std::vector<std::string>* errors);

 private:
  bool ReconcileOne(std::string profile, const std::string objs[8], ProfileStats& st, std::string* error);
  Store* store_;
  nbrm_ctx* ctx_;
  std::string wi_, labels_;
  PluginFn apply_, revoke_;
};

}  // namespace nbrm_host
