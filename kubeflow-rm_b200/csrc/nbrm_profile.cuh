// nbrm_profile.cuh — the objects one Profile reconcile builds.
//
// Reference: components/profile-controller/controllers/profile_controller.go —
// Reconcile :105-334: Namespace :127-137 (labels through setNamespaceLabels
// :754-772, rendered once on the host from the labels file), Istio
// AuthorizationPolicy getAuthorizationPolicy :419-504 / metadata :512-520,
// ServiceAccounts and their RoleBindings updateServiceAccount :592-637, the owner
// RoleBinding :230-246, ResourceQuota :253-262; every object gets the
// controller ownerReference to the Profile (SetControllerReference :138,:523,
// :604,:645).  Input slot 0: Profile CRs.  Output: 8 objects per Profile in the
// order Namespace, AuthorizationPolicy, SA default-editor, RoleBinding
// default-editor, SA default-viewer, RoleBinding default-viewer, RoleBinding
// namespaceAdmin, ResourceQuota (length 0 without spec.resourceQuotaSpec.hard).
#pragma once
#include "nbrm_nbc.cuh"

namespace nbrm {

struct ProfView {
  Span name, uid, owner, owner_name, quota;
  bool has_hard;
};

NB_HD ProfView view_profile(const Doc& d) {
  ProfView v;
  v.name = v.uid = v.owner = v.owner_name = v.quota = mk(0, 0);
  v.has_hard = false;
  const Span rt = root(d);
  const Span meta = FIND(d, rt, "metadata");
  v.name = FIND(d, meta, "name");
  v.uid = FIND(d, meta, "uid");
  const Span spec = FIND(d, rt, "spec");
  v.owner = FIND(d, spec, "owner");
  v.owner_name = FIND(d, v.owner, "name");
  v.quota = FIND(d, spec, "resourceQuotaSpec");
  const Span hard = FIND(d, v.quota, "hard");
  v.has_hard = is_object(d, hard) && !is_empty_container(d, hard);  // len(Hard) > 0 (:253)
  if (!is_string(d, v.name)) v.name = mk(0, 0);
  if (!is_string(d, v.uid)) v.uid = mk(0, 0);
  if (!is_string(d, v.owner_name)) v.owner_name = mk(0, 0);
  return v;
}

NB_HD void plan_prof_owner(Planner& P, const ProfView& v) {
  LIT(P, PROF_OWNER);
  put_str(P, v.name);
  LIT(P, OWNER_UID);
  put_str(P, v.uid);
  LIT(P, OWNER_TAIL);
}

NB_HD void plan_profile_object(Planner& P, const Doc& d, const ProfView& v, const EnvLits& env,
                               uint32_t k) {
  switch (k) {
    case 0:  // Namespace
      LIT(P, PROF_NS_HEAD);
      put_str(P, v.name);
      LIT(P, PROF_NS_ANN);
      put_str(P, v.owner_name);
      LIT(P, PROF_NS_LABELS);
      P.pool(env.prof_ns_labels.off, env.prof_ns_labels.len);
      LIT(P, RBRACE);
      plan_prof_owner(P, v);
      LIT(P, RBRACE2);
      break;
    case 1:  // AuthorizationPolicy ns-owner-access-istio
      LIT(P, PROF_AUTHZ_HEAD);
      put_str(P, v.name);
      LIT(P, PROF_ANN_USER);
      put_str(P, v.owner_name);
      LIT(P, PROF_ANN_ROLE);
      plan_prof_owner(P, v);
      LIT(P, PROF_AUTHZ_RULE1);
      P.pool(env.prof_igw.off, env.prof_igw.len);
      LIT(P, COMMA);
      P.pool(env.prof_kfp.off, env.prof_kfp.len);
      LIT(P, PROF_AUTHZ_WHEN);
      P.pool(env.prof_when_key.off, env.prof_when_key.len);
      LIT(P, PROF_AUTHZ_VALUES);
      P.pool(env.prof_prefix.off, env.prof_prefix.len);
      P.in(inner(v.owner_name));
      LIT(P, PROF_AUTHZ_RULE2);
      put_str(P, v.name);
      LIT(P, PROF_AUTHZ_RULE34);
      P.pool(env.prof_nbc.off, env.prof_nbc.len);
      LIT(P, PROF_AUTHZ_TAIL);
      break;
    case 2:
    case 4:  // ServiceAccounts
      if (k == 2) LIT(P, PROF_SA_EDITOR);
      else LIT(P, PROF_SA_VIEWER);
      put_str(P, v.name);
      plan_prof_owner(P, v);
      LIT(P, RBRACE2);
      break;
    case 3:
    case 5:  // their RoleBindings
      if (k == 3) LIT(P, PROF_RB_EDITOR);
      else LIT(P, PROF_RB_VIEWER);
      put_str(P, v.name);
      plan_prof_owner(P, v);
      if (k == 3) LIT(P, PROF_RB_EDITOR_REF);
      else LIT(P, PROF_RB_VIEWER_REF);
      put_str(P, v.name);
      LIT(P, PROF_RB_SUBJ_TAIL);
      break;
    case 6:  // owner RoleBinding namespaceAdmin
      LIT(P, PROF_RB_ADMIN);
      put_str(P, v.name);
      LIT(P, PROF_ANN_USER);
      put_str(P, v.owner_name);
      LIT(P, PROF_ANN_ROLE);
      plan_prof_owner(P, v);
      LIT(P, PROF_RB_ADMIN_REF);
      if (is_object(d, v.owner)) P.in(v.owner);
      else {
        LIT(P, LBRACE);
        LIT(P, RBRACE);
      }
      LIT(P, PROF_RB_ADMIN_TAIL);
      break;
    case 7:  // ResourceQuota kf-resource-quota
      if (!v.has_hard) break;
      LIT(P, PROF_QUOTA_HEAD);
      put_str(P, v.name);
      plan_prof_owner(P, v);
      LIT(P, PROF_QUOTA_SPEC);
      P.in(v.quota);
      LIT(P, RBRACE);
      break;
  }
}

struct GenProfile {
  static const uint32_t NOBJ = 8;
  static const bool SHARED1 = false;
  static const uint32_t CAP = 32;
  struct K {
    DocCtx& c;
    const ProfView& v;
    NB_HD void operator()(Planner& P, uint32_t k) { plan_profile_object(P, c.d[0], v, *c.env, k); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    const Doc& d = c.d[0];
    const uint32_t i0 = skip_ws(d, 0);
    const uint32_t st = (i0 < d.n && d.at(i0) == '{') ? E_OK : E_PARSE;
    ProfView v = view_profile(d);
    K k{c, v};
    return plan_objects(c, s, NOBJ, st, k);
  }
};

}  // namespace nbrm
