// nbrm_convert.cuh — Notebook version conversion.
//
// Reference: components/notebook-controller/api/v1/notebook_conversion.go:25-69 (ConvertTo,
// ConvertFrom; api/v1alpha1/notebook_conversion.go holds the same pair; the hub is v1beta1).
// Both directions copy exactly: spec.template.spec, status.readyReplicas,
// status.containerState, and of each condition type, lastProbeTime, reason, message.  A
// condition's status and lastTransitionTime are not copied, ObjectMeta is not touched.
//
// GenConvert: slot 0 = Notebooks of any version.  arg0 = the version to convert to
// (0 kubeflow.org/v1beta1, 1 kubeflow.org/v1, 2 kubeflow.org/v1alpha1).  flags bit 0
// (CONVERT_KEEP_METADATA): carry metadata through (the reference's functions leave the
// destination's ObjectMeta as it was — empty for a fresh object).
#pragma once
#include "nbrm_status.cuh"

namespace nbrm {

enum ConvertFlags : uint32_t { CONVERT_KEEP_METADATA = 1u };

NB_HD void plan_convert(Planner& P, DocCtx& c) {
  const Doc& d = c.d[0];
  Span meta = mk(0, 0), podspec = mk(0, 0), status = mk(0, 0);
  {
    ObjIter it(d, root(d));
    Member m;
    while (it.next(d, m)) {
      if (KEY_IS(d, m, "metadata")) meta = m.val;
      else if (KEY_IS(d, m, "status")) status = m.val;
      else if (KEY_IS(d, m, "spec")) podspec = FIND(d, FIND(d, m.val, "template"), "spec");
    }
  }
  if (c.arg0 == 1) LIT(P, CONVERT_HEAD_V1);
  else if (c.arg0 == 2) LIT(P, CONVERT_HEAD_V1ALPHA1);
  else LIT(P, CONVERT_HEAD_V1BETA1);  // {"apiVersion":"...","kind":"Notebook","metadata":
  if ((c.flags & CONVERT_KEEP_METADATA) && is_object(d, meta)) P.in(meta);
  else {
    LIT(P, LBRACE);
    LIT(P, RBRACE);
  }
  LIT(P, CONVERT_SPEC);  // ,"spec":{"template":{"spec":
  if (is_object(d, podspec)) P.in(podspec);
  else {
    LIT(P, LBRACE);
    LIT(P, RBRACE);
  }
  LIT(P, CONVERT_STATUS);  // }},"status":{"conditions":[
  Sep cs;
  ArrIter it(d, FIND(d, status, "conditions"));
  Span e;
  while (it.next(d, e)) {
    Span ty = mk(0, 0), lp = mk(0, 0), rs = mk(0, 0), ms = mk(0, 0);
    ObjIter ci(d, e);
    Member m;
    while (ci.next(d, m)) {
      if (KEY_IS(d, m, "type")) ty = m.val;
      else if (KEY_IS(d, m, "lastProbeTime")) lp = m.val;
      else if (KEY_IS(d, m, "reason")) rs = m.val;
      else if (KEY_IS(d, m, "message")) ms = m.val;
    }
    cs.next(P);
    LIT(P, COND_TYPE);
    if (nonempty_string(d, ty)) P.in(ty); else LIT(P, EMPTY_STR);
    LIT(P, COND_STATUS);
    LIT(P, EMPTY_STR);  // not copied (:31-36)
    if (nonempty_string(d, lp)) {
      LIT(P, COND_PROBE);
      P.in(lp);
    }
    if (nonempty_string(d, rs)) {
      LIT(P, COND_REASON);
      P.in(rs);
    }
    if (nonempty_string(d, ms)) {
      LIT(P, COND_MESSAGE);
      P.in(ms);
    }
    LIT(P, RBRACE);
  }
  LIT(P, STATUS_READY);
  const Span rr = FIND(d, status, "readyReplicas");
  if (rr.ok() && !is_null(d, rr)) P.in(rr); else LIT(P, ZERO);
  LIT(P, STATUS_STATE);
  const Span st = FIND(d, status, "containerState");
  if (is_object(d, st)) P.in(st);
  else {
    LIT(P, LBRACE);
    LIT(P, RBRACE);
  }
  LIT(P, RBRACE);
  LIT(P, RBRACE);
}

struct GenConvert {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;
  static const uint32_t CAP = 160;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { plan_convert(P, c); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    K k{c};
    const uint32_t st = is_object(c.d[0], root(c.d[0])) ? E_OK : E_PARSE;
    return plan_objects(c, s, NOBJ, st, k);
  }
};

}  // namespace nbrm
