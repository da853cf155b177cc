// nbrm_status.cuh — Notebook status mirroring.
//
// Reference: createNotebookStatus, components/notebook-controller/controllers/
// notebook_controller.go:290-349, and PodCondToNotebookCond :351-390; golden
// cases notebook_controller_test.go:103-275.
// Input slot 0: Notebook, slot 1: its StatefulSet, slot 2: its Pod "<name>-0"
// (empty documents = not found).  Output: the NotebookStatus object.
#pragma once
#include "nbrm_cmp.cuh"
#include "nbrm_time.cuh"

namespace nbrm {

// reflect.DeepEqual(pod.Status, corev1.PodStatus{}) (:304): every member is a
// zero scalar; any slice or struct value, even an empty one, makes it non-zero.
NB_HD bool pod_status_is_zero(const Doc& d, Span st) {
  if (!is_object(d, st)) return true;
  ObjIter it(d, st);
  Member m;
  while (it.next(d, m)) {
    const Span v = m.val;
    if (is_null(d, v)) continue;
    const char c = d.at(v.off);
    if (c == '"') {
      if (v.len > 2) return false;
    } else if (c == '{' || c == '[') {
      return false;
    } else if (!(v.len == 1 && c == '0') && !(v.len == 5 && c == 'f')) {
      return false;
    }
  }
  return true;
}

// corev1.ContainerState with all three pointers nil.
NB_HD bool container_state_is_zero(const Doc& d, Span st) {
  if (!is_object(d, st)) return true;
  return is_nil(d, FIND(d, st, "waiting")) && is_nil(d, FIND(d, st, "running")) &&
         is_nil(d, FIND(d, st, "terminated"));
}

NB_HD bool nonempty_string(const Doc& d, Span v) { return is_string(d, v) && v.len > 2; }

NB_HD void plan_status(Planner& P, DocCtx& c) {
  const Doc& nb = c.d[0];
  const Doc& sts = c.d[1];
  const Doc& pod = c.d[2];
  LIT(P, STATUS_HEAD);  // {"conditions":[
  // each document is descended once: only a few members of each matter here
  Span pst = mk(0, 0), nbname = mk(0, 0), cur = mk(0, 0), rr = mk(0, 0);
  if (pod.n) {
    Cursor pc;
    pc.i = 0;
    Member m;
    if (obj_enter(pod, pc))
      while (obj_key(pod, pc, m)) {
        const Span val = val_skip(pod, pc);
        if (KEY_IS(pod, m, "status")) pst = val;
      }
  }
  const ViewIn& nv = c.view[0];
  const Span vroot = nv.ok() ? nv.get(VF_ROOT) : mk(0, 0);
  const Span vmeta = nv.ok() ? nv.get(VF_META) : mk(0, 0), vspec = nv.ok() ? nv.get(VF_SPEC) : mk(0, 0);
  if (vroot.ok() && vmeta.ok() && vspec.ok() && vroot.end() <= nb.n) {
    // the generator that wrote this Notebook left the extents of its metadata and spec: neither is
    // skipped through, and inside metadata neither are the labels and annotations
    const Span vl = nv.get(VF_LABELS), va = nv.get(VF_ANNOTATIONS);
    ObjIter it(nb, vroot);
    Member m;
    while (it.next(nb, m, vmeta, vspec)) {
      if (KEY_IS(nb, m, "metadata")) {
        ObjIter mi(nb, m.val);
        Member mm;
        while (mi.next(nb, mm, vl, va))
          if (KEY_IS(nb, mm, "name")) nbname = mm.val;
      } else if (KEY_IS(nb, m, "status")) {
        ObjIter si(nb, m.val);
        Member mm;
        while (si.next(nb, mm))
          if (KEY_IS(nb, mm, "containerState")) cur = mm.val;
      }
    }
  } else {
    Cursor nc;
    nc.i = 0;
    Member m;
    if (obj_enter(nb, nc))
      while (obj_key(nb, nc, m)) {
        if (KEY_IS(nb, m, "metadata") && obj_enter(nb, nc)) {
          Member mm;
          while (obj_key(nb, nc, mm)) {
            const Span val = val_skip(nb, nc);
            if (KEY_IS(nb, mm, "name")) nbname = val;
          }
        } else if (KEY_IS(nb, m, "status") && obj_enter(nb, nc)) {
          Member mm;
          while (obj_key(nb, nc, mm)) {
            const Span val = val_skip(nb, nc);
            if (KEY_IS(nb, mm, "containerState")) cur = val;
          }
        } else {
          val_skip(nb, nc);
        }
      }
  }
  if (sts.n) {
    Cursor sc;
    sc.i = 0;
    Member m;
    if (obj_enter(sts, sc))
      while (obj_key(sts, sc, m)) {
        if (KEY_IS(sts, m, "status") && obj_enter(sts, sc)) {
          Member mm;
          while (obj_key(sts, sc, mm)) {
            const Span val = val_skip(sts, sc);
            if (KEY_IS(sts, mm, "readyReplicas")) rr = val;
          }
        } else {
          val_skip(sts, sc);
        }
      }
  }
  const bool have_pod = !pod_status_is_zero(pod, pst);
  Span state = mk(0, 0);
  bool state_set = false;
  if (have_pod) {
    // "now" as metav1.Now() would stamp it (:377,:386), once per Notebook
    uint32_t now_off = 0, now_len = 0;
    Sep cs;
    ArrIter it(pod, FIND(pod, pst, "conditions"));
    Span e;
    while (it.next(pod, e)) {  // :341-346
      Span ty = mk(0, 0), stv = mk(0, 0), lp = mk(0, 0), lt = mk(0, 0), rs = mk(0, 0), ms = mk(0, 0);
      ObjIter ci(pod, e);
      Member m;
      while (ci.next(pod, m)) {
        if (KEY_IS(pod, m, "type")) ty = m.val;
        else if (KEY_IS(pod, m, "status")) stv = m.val;
        else if (KEY_IS(pod, m, "lastProbeTime")) lp = m.val;
        else if (KEY_IS(pod, m, "lastTransitionTime")) lt = m.val;
        else if (KEY_IS(pod, m, "reason")) rs = m.val;
        else if (KEY_IS(pod, m, "message")) ms = m.val;
      }
      if ((!nonempty_string(pod, lp) || !nonempty_string(pod, lt)) && !now_len) {
        if (c.aux.used + 32 <= c.aux.cap) {
          char* o = c.aux.p + c.aux.used;
          o[0] = '"';
          now_len = format_rfc3339((long long)c.now, 0, o + 1) + 2;
          o[now_len - 1] = '"';
          now_off = c.aux.abs + c.aux.used;
          c.aux.used += now_len;
        } else {
          P.err = E_LIMIT;
          return;
        }
      }
      cs.next(P);
      LIT(P, COND_TYPE);  // {"type":
      if (nonempty_string(pod, ty)) P.in(ty, 2); else LIT(P, EMPTY_STR);
      LIT(P, COND_STATUS);
      if (nonempty_string(pod, stv)) P.in(stv, 2); else LIT(P, EMPTY_STR);
      LIT(P, COND_PROBE);
      if (nonempty_string(pod, lp)) P.in(lp, 2); else P.aux(now_off, now_len);
      LIT(P, COND_TRANSITION);
      if (nonempty_string(pod, lt)) P.in(lt, 2); else P.aux(now_off, now_len);
      if (nonempty_string(pod, rs)) {
        LIT(P, COND_REASON);
        P.in(rs, 2);
      }
      if (nonempty_string(pod, ms)) {
        LIT(P, COND_MESSAGE);
        P.in(ms, 2);
      }
      LIT(P, RBRACE);
    }
    // :314-330 the state of the container named like the Notebook
    const bool cur_zero = container_state_is_zero(nb, cur);
    ArrIter si(pod, FIND(pod, pst, "containerStatuses"));
    while (si.next(pod, e)) {
      const Span nm = FIND(pod, e, "name");
      const bool same = is_string(nb, nbname) ? (is_string(pod, nm) && str_equal(pod, nm, nb, nbname))
                                              : (!is_string(pod, nm) || nm.len == 2);
      if (!same) continue;
      const Span st = FIND(pod, e, "state");
      if (container_state_is_zero(pod, st) && cur_zero) continue;  // Go == on three nil pointers (:319)
      state = st;
      state_set = true;
      break;
    }
  }
  LIT(P, STATUS_READY);  // ],"readyReplicas":
  if (rr.ok() && !is_null(sts, rr)) P.in(rr, 1); else LIT(P, ZERO);
  LIT(P, STATUS_STATE);  // ,"containerState":
  if (state_set && is_object(pod, state)) P.in(state, 2);
  else {
    LIT(P, LBRACE);
    LIT(P, RBRACE);
  }
  LIT(P, RBRACE);
}

struct GenStatus {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;  // slot 1 is per-tuple, not one shared arena
  static const uint32_t CAP = 160;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { plan_status(P, c); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    K k{c};
    const Doc& nb = c.d[0];
    const uint32_t i0 = skip_ws(nb, 0);
    uint32_t st = (i0 < nb.n && nb.at(i0) == '{') ? E_OK : E_PARSE;
    return plan_objects(c, s, NOBJ, st, k);
  }
};

}  // namespace nbrm
