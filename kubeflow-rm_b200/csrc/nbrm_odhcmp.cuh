// nbrm_odhcmp.cuh — what the ODH reconcilers write for one (desired, found) pair.
//
// Reference: components/odh-notebook-controller/controllers —
//   reconcileRoute            notebook_route.go:76-139, CompareNotebookRoutes :65-72
//   reconcileNetworkPolicy    notebook_network.go:66-122, CompareNotebookNetworkPolicies :124-128
//   ReconcileOAuthServiceAccount / Service / Secret   notebook_oauth.go:77-113, :151-187, :214-250
//                             (Get, Create when NotFound, nothing otherwise)
//   reconcileRoleBinding      notebook_rbac.go:89-140 (Create when NotFound; Update(desired)
//                             when the subjects differ)
//
// Input slot 0 holds the desired objects as GenOdh emits them (an empty
// document = the reconcile builds no such object), slot 1 the found objects (an
// empty document = NotFound); tuple i pairs document i of both.  Output per pair:
//   found missing              -> the desired object        result 1 (caller Creates it)
//   Route / NetworkPolicy whose labels or spec differ
//                              -> the found object with spec and metadata.labels
//                                 replaced by the desired ones   result 2 (caller Updates it)
//   RoleBinding whose subjects differ
//                              -> the desired object without ownerReferences (the reference
//                                 only sets them on the Create branch)     result 2
//   anything else              -> length 0                  result 0
#pragma once
#include "nbrm_cmp.cuh"

namespace nbrm {

enum OdhWrite : uint32_t { ODH_NONE = 0, ODH_CREATE = 1, ODH_UPDATE = 2 };

// DeepEqual of two Route specs with spec.host left out on both sides (:67).
NB_HD bool route_spec_equal(const Doc& a, Span x, const Doc& b, Span y, bool* deep) {
  const bool xo = is_object(a, x), yo = is_object(b, y);
  if (!xo || !yo) return json_equal(a, x, b, y, deep);
  uint32_t na = 0, nb = 0;
  {
    ObjIter it(b, y);
    Member m;
    while (it.next(b, m))
      if (!KEY_IS(b, m, "host") && !is_nil(b, m.val)) ++nb;
  }
  ObjIter it(a, x);
  Member m;
  while (it.next(a, m)) {
    if (KEY_IS(a, m, "host") || is_nil(a, m.val)) continue;
    ++na;
    const Span bv = find_key(b, y, a, m.key);
    if (is_nil(b, bv) || !json_equal(a, m.val, b, bv, deep)) return false;
  }
  return na == nb;
}

// found with spec and metadata.labels taken from desired (route.go:128-129, network.go:111-112)
NB_HD void plan_spec_labels_update(Planner& P, const Doc& from, Span fspec, Span flabels,
                                   const Doc& to) {
  LIT(P, LBRACE);
  Sep s;
  bool had_spec = false;
  ObjIter it(to, root(to));
  Member m;
  while (it.next(to, m)) {
    if (KEY_IS(to, m, "metadata") && is_object(to, m.val)) {
      s.next(P);
      LIT(P, KEY_METADATA);
      LIT(P, LBRACE);
      Sep ms;
      ObjIter mi(to, m.val);
      Member mm;
      while (mi.next(to, mm)) {
        if (KEY_IS(to, mm, "labels")) continue;
        ms.in(P, to, mm.all, 1);
      }
      if (!is_nil(from, flabels)) {
        ms.next(P);
        LIT(P, KEY_LABELS);
        P.in(flabels, 0);
      }
      LIT(P, RBRACE);
    } else if (KEY_IS(to, m, "spec")) {
      had_spec = true;
      if (!is_nil(from, fspec)) {
        s.next(P);
        LIT(P, KEY_SPEC);
        P.in(fspec, 0);
      }
    } else {
      s.in(P, to, m.all, 1);
    }
  }
  if (!had_spec && !is_nil(from, fspec)) {
    s.next(P);
    LIT(P, KEY_SPEC);
    P.in(fspec, 0);
  }
  LIT(P, RBRACE);
}

// desired as it is, minus metadata.ownerReferences (rbac.go:104,133)
NB_HD void plan_without_owner(Planner& P, const Doc& from) {
  LIT(P, LBRACE);
  Sep s;
  ObjIter it(from, root(from));
  Member m;
  while (it.next(from, m)) {
    if (KEY_IS(from, m, "metadata") && is_object(from, m.val)) {
      s.next(P);
      LIT(P, KEY_METADATA);
      LIT(P, LBRACE);
      Sep ms;
      ObjIter mi(from, m.val);
      Member mm;
      while (mi.next(from, mm)) {
        if (KEY_IS(from, mm, "ownerReferences")) continue;
        ms.in(P, from, mm.all, 0);
      }
      LIT(P, RBRACE);
    } else {
      s.in(P, from, m.all, 0);
    }
  }
  LIT(P, RBRACE);
}

NB_HD uint32_t plan_odh_decide(Planner& P, const Doc& from, const Doc& to, const ViewIn& fview) {
  const Span fr = fview.ok() ? fview.get(VF_ROOT) : root(from);
  if (!is_object(from, fr)) {
    P.err = E_PARSE;
    return ODH_NONE;
  }
  if (to.n == 0 || to.at(skip_ws(to, 0)) != '{') {  // NotFound -> Create(desired)
    P.in(fr, 0);
    return ODH_CREATE;
  }
  // one walk over each root: kind, metadata, spec, subjects
  Span kind = mk(0, 0), fmeta = mk(0, 0), fspec = mk(0, 0), fsubj = mk(0, 0), flabels = mk(0, 0);
  if (fview.ok()) {  // the generator of the desired object wrote down where these lie (nbrm_plan.cuh, views)
    kind = fview.get(VF_KIND);
    if (!str_eq(from, kind, "Route") && !str_eq(from, kind, "NetworkPolicy") && !str_eq(from, kind, "RoleBinding"))
      return ODH_NONE;
    fspec = fview.get(VF_SPEC);
    fsubj = fview.get(VF_SUBJECTS);
    flabels = fview.get(VF_LABELS);
  } else {
    ObjIter it(from, fr);
    Member m;
    while (it.next(from, m)) {
      if (KEY_IS(from, m, "kind")) {
        kind = m.val;
        // ServiceAccount, Service, Secret: created once, never compared (kind leads the object)
        if (!str_eq(from, kind, "Route") && !str_eq(from, kind, "NetworkPolicy") &&
            !str_eq(from, kind, "RoleBinding"))
          return ODH_NONE;
      } else if (KEY_IS(from, m, "metadata")) fmeta = m.val;
      else if (KEY_IS(from, m, "spec")) fspec = m.val;
      else if (KEY_IS(from, m, "subjects")) fsubj = m.val;
    }
    flabels = FIND(from, fmeta, "labels");
  }
  const bool is_route = str_eq(from, kind, "Route");
  const bool is_np = !is_route && str_eq(from, kind, "NetworkPolicy");
  const bool is_rb = !is_route && !is_np && str_eq(from, kind, "RoleBinding");
  if (!is_route && !is_np && !is_rb) return ODH_NONE;  // no kind at all
  Span tmeta = mk(0, 0), tspec = mk(0, 0), tsubj = mk(0, 0);
  {
    ObjIter it(to, root(to));
    Member m;
    while (it.next(to, m)) {
      if (KEY_IS(to, m, "metadata")) tmeta = m.val;
      else if (KEY_IS(to, m, "spec")) tspec = m.val;
      else if (KEY_IS(to, m, "subjects")) tsubj = m.val;
    }
  }
  bool deep = false;
  if (is_rb) {
    if (json_equal(from, fsubj, to, tsubj, &deep)) return ODH_NONE;
    if (deep) {
      P.err = E_LIMIT;
      return ODH_NONE;
    }
    plan_without_owner(P, from);
    return ODH_UPDATE;
  }
  const Span tlabels = FIND(to, tmeta, "labels");
  bool same = json_equal(from, flabels, to, tlabels, &deep);
  if (same) same = is_route ? route_spec_equal(from, fspec, to, tspec, &deep) : json_equal(from, fspec, to, tspec, &deep);
  if (deep) {
    P.err = E_LIMIT;  // nested deeper than the compare follows: no answer, not "differs"
    return ODH_NONE;
  }
  if (same) return ODH_NONE;
  plan_spec_labels_update(P, from, fspec, flabels, to);
  return ODH_UPDATE;
}

// The ODH write decision: slot 0 desired, slot 1 found, one output per pair.
struct GenOdhDecide {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;
  static const uint32_t CAP = 48;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { c.result = plan_odh_decide(P, c.d[0], c.d[1], c.view[0]); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    K k{c};
    uint32_t st = E_OK;
    if (c.d[0].n == 0) {  // the reconcile builds no such object for this Notebook
      s.nseg[0] = 0;
      s.lens[0] = 0;
      return E_OK;
    }
    const uint32_t r = plan_objects(c, s, NOBJ, st, k);
    if (r) c.result = ODH_NONE;
    return r;
  }
};

}  // namespace nbrm
