// nbrm_cull.cuh — one idle-culling pass per Notebook.
//
// Reference: CullingReconciler.Reconcile and its helpers, components/
// notebook-controller/controllers/culling_controller.go:85-169 (the pass),
// :173-183 cullingCheckPeriodHasPassed, :186-207 notebookIsIdle, :281-293
// allKernelsAreIdle, :295-315 getNotebookRecentTime, :337-355
// compareAnnotationTimeToResource, :357-414 updateTimestampFrom{Kernels,
// Terminals}Activity, :416-458 annotation bookkeeping, :461-488 stop annotation.
// Golden: culling_controller_test.go:16-249.
//
// Input slot 0: Notebook; slot 1: body of GET /api/kernels (empty document =
// the GET failed); slot 2: body of GET /api/terminals.  extra: one byte per
// Notebook, non-zero when the Pod "<name>-0" exists.  arg0 = CULL_IDLE_TIME,
// arg1 = IDLENESS_CHECK_PERIOD (minutes), now = time.Now() in Unix seconds.
// Output: the Notebook to Update (length 0 when the pass changes nothing), and
// result = outcome | updates << 8 with the outcome codes below.
#pragma once
#include "nbrm_gen.cuh"
#include "nbrm_time.cuh"

namespace nbrm {

enum CullOutcome : uint32_t {
  CULL_REMOVED_STOPPED = 1,  // stop annotation set: culling annotations removed (:103-111)
  CULL_REMOVED_NO_POD = 2,   // pod missing: culling annotations removed (:116-125)
  CULL_WAIT = 3,             // check period not over: requeue (:141-144)
  CULL_CHECKED = 4,          // annotations refreshed, not idle (:147-153,:168)
  CULL_CULLED = 5,           // idle: stop annotation set (:156-167)
};

// One managed annotation's final value.
struct AnnVal {
  uint32_t kind;  // 0 absent, 1 the original value (span in slot 0), 2 bytes in aux
  Span raw;
  uint32_t aux_off, aux_len;  // with quotes
  Time t;                     // parsed value (ok == false when it does not parse)
};

NB_HD AnnVal ann_from_span(const Doc& d, Span v) {
  AnnVal a;
  a.kind = v.ok() ? 1u : 0u;
  a.raw = v;
  a.aux_off = a.aux_len = 0;
  a.t.ok = false;
  if (is_string(d, v)) a.t = parse_rfc3339(d, inner(v));
  return a;
}

// Write "<RFC 3339 time>" (quoted) into aux.
NB_HD bool ann_set_time(AnnVal& a, Aux& aux, long long secs, int off) {
  if (aux.used + 32 > aux.cap) return false;
  char* o = aux.p + aux.used;
  o[0] = '"';
  const uint32_t n = format_rfc3339(secs, off, o + 1);
  o[n + 1] = '"';
  a.kind = 2;
  a.aux_off = aux.abs + aux.used;
  a.aux_len = n + 2;
  aux.used += n + 2;
  a.t.secs = secs;
  a.t.nanos = 0;
  a.t.off = off;
  a.t.ok = true;
  return true;
}

// getNotebookRecentTime over the "last_activity" of every element (:295-315),
// with allKernelsAreIdle (:281-293) when check_idle.  Returns false when the
// list is nil or empty.  all_idle / ok / recent are outputs.
NB_HD bool scan_activity(const Doc& d, bool check_idle, bool& all_idle, bool& ok, Time& recent) {
  all_idle = true;
  ok = true;
  recent.ok = false;
  if (!d.n) return false;
  const Span arr = root(d);
  if (!is_array(d, arr)) return false;
  uint32_t n = 0;
  ArrIter it(d, arr);
  Span e;
  while (it.next(d, e)) {
    Span la = mk(0, 0), es = mk(0, 0);
    ObjIter oi(d, e);
    Member m;
    while (oi.next(d, m)) {
      if (KEY_IS(d, m, "last_activity")) la = m.val;
      else if (KEY_IS(d, m, "execution_state")) es = m.val;
    }
    if (check_idle && !str_eq(d, es, "idle")) all_idle = false;
    Time t;
    t.ok = false;
    if (is_string(d, la)) t = parse_rfc3339(d, inner(la));
    if (!t.ok) ok = false;
    else if (n == 0 || (recent.ok && time_after(t, recent))) recent = t;
    ++n;
  }
  return n > 0;
}

NB_HD void plan_cull(Planner& P, DocCtx& c) {
  const Doc& nb = c.d[0];
  const long long now = (long long)c.now;
  const long long idle_s = (long long)(int)c.arg0 * 60, period_s = (long long)(int)c.arg1 * 60;
  const Span rootv = root(nb);
  Member meta_m;
  meta_m.val = mk(0, 0);
  Span ann = mk(0, 0);
  Span v_stop = mk(0, 0), v_la = mk(0, 0), v_chk = mk(0, 0);
  {
    ObjIter it(nb, rootv);
    Member m;
    while (it.next(nb, m))
      if (KEY_IS(nb, m, "metadata")) meta_m = m;
    ann = FIND(nb, meta_m.val, "annotations");
    ObjIter ai(nb, ann);
    while (ai.next(nb, m)) {
      if (KEY_IS(nb, m, "kubeflow-resource-stopped")) v_stop = m.val;
      else if (KEY_IS(nb, m, "notebooks.kubeflow.org/last-activity")) v_la = m.val;
      else if (KEY_IS(nb, m, "notebooks.kubeflow.org/last_activity_check_timestamp")) v_chk = m.val;
    }
  }
  AnnVal stop = ann_from_span(nb, v_stop), la = ann_from_span(nb, v_la), chk = ann_from_span(nb, v_chk);
  const bool pod_exists = c.extra ? ((const unsigned char*)c.extra)[c.index] != 0 : true;
  uint32_t outcome, updates = 0;
  if (stop.kind) {  // :103-111
    la.kind = chk.kind = 0;
    outcome = CULL_REMOVED_STOPPED;
    updates = 1;
  } else if (!pod_exists) {  // :116-125
    la.kind = chk.kind = 0;
    outcome = CULL_REMOVED_NO_POD;
    updates = 1;
  } else {
    bool okw = true;
    if (!(la.kind && chk.kind)) {  // :131-138 initializeAnnotations
      okw &= ann_set_time(la, c.aux, now, 0);
      chk = la;
      updates = 1;
    }
    // :173-183, a stored stamp that does not parse reads as the zero time (:178)
    // `now` is whole seconds, so stored + period < now holds exactly when the
    // whole-second parts say so, whatever fraction the stored stamp carries
    const long long stored = chk.t.ok ? chk.t.secs : -62135596800ll;
    const bool passed = stored + period_s < now;
    if (!passed) {
      outcome = CULL_WAIT;
    } else {
      // :318-335 updateNotebookLastActivityAnnotation
      bool idle_all, okt;
      Time recent;
      if (scan_activity(c.d[1], true, idle_all, okt, recent)) {  // :357-388
        if (!idle_all) {
          okw &= ann_set_time(la, c.aux, now, 0);  // a busy kernel: last-activity = now (:364-372)
        } else if (okt && la.t.ok && !time_after(la.t, recent)) {  // :379-385, never backwards (:350)
          okw &= ann_set_time(la, c.aux, recent.secs, recent.off);
        }
      }
      if (scan_activity(c.d[2], false, idle_all, okt, recent)) {  // :390-414
        if (okt && la.t.ok && !time_after(la.t, recent))
          okw &= ann_set_time(la, c.aux, recent.secs, recent.off);
      }
      okw &= ann_set_time(chk, c.aux, now, 0);  // :416-424
      ++updates;
      // :186-207 notebookIsIdle
      // time.Now().After(lastActivity + CULL_IDLE_TIME): with `now` in whole
      // seconds that is now > the whole-second part of the cap
      const bool idle = la.t.ok && now > la.t.secs + idle_s;
      if (idle) {
        okw &= ann_set_time(stop, c.aux, now, 0);  // :461-478
        ++updates;
        outcome = CULL_CULLED;
      } else {
        outcome = CULL_CHECKED;
      }
    }
    if (!okw) {
      P.err = E_LIMIT;
      return;
    }
  }
  c.result = outcome | (updates << 8);
  if (!updates) return;  // nothing to write back

  // the Notebook with its managed annotations replaced
  const bool any_final = stop.kind || la.kind || chk.kind;
  LIT(P, LBRACE);
  Sep s;
  bool had_meta = false;
  ObjIter it(nb, rootv);
  Member m;
  while (it.next(nb, m)) {
    if (!KEY_IS(nb, m, "metadata") || !is_object(nb, m.val)) {
      s.in(P, nb, m.all);
      continue;
    }
    had_meta = true;
    s.next(P);
    LIT(P, KEY_METADATA);
    LIT(P, LBRACE);
    Sep ms;
    bool had_ann = false;
    ObjIter mi(nb, m.val);
    Member mm;
    while (mi.next(nb, mm)) {
      if (!KEY_IS(nb, mm, "annotations")) {
        ms.in(P, nb, mm.all);
        continue;
      }
      had_ann = true;
      if (!is_object(nb, mm.val) && !any_final) {  // e.g. null: nothing to add, leave as is
        ms.in(P, nb, mm.all);
        continue;
      }
      ms.next(P);
      LIT(P, KEY_ANNOTATIONS);
      LIT(P, LBRACE);
      Sep as;
      ObjIter ai(nb, mm.val);
      Member am;
      while (ai.next(nb, am)) {
        if (KEY_IS(nb, am, "kubeflow-resource-stopped") ||
            KEY_IS(nb, am, "notebooks.kubeflow.org/last-activity") ||
            KEY_IS(nb, am, "notebooks.kubeflow.org/last_activity_check_timestamp"))
          continue;
        as.in(P, nb, am.all);
      }
      const AnnVal* fin[3] = {&stop, &la, &chk};
      for (int k = 0; k < 3; ++k) {
        if (!fin[k]->kind) continue;
        as.next(P);
        if (k == 0) LIT(P, ANN_STOP);
        else if (k == 1) LIT(P, ANN_LAST_ACTIVITY);
        else LIT(P, ANN_LAST_CHECK);
        if (fin[k]->kind == 1) P.in(fin[k]->raw);
        else P.aux(fin[k]->aux_off, fin[k]->aux_len);
      }
      LIT(P, RBRACE);
    }
    if (!had_ann && any_final) {
      ms.next(P);
      LIT(P, KEY_ANNOTATIONS);
      LIT(P, LBRACE);
      Sep as;
      const AnnVal* fin[3] = {&stop, &la, &chk};
      for (int k = 0; k < 3; ++k) {
        if (!fin[k]->kind) continue;
        as.next(P);
        if (k == 0) LIT(P, ANN_STOP);
        else if (k == 1) LIT(P, ANN_LAST_ACTIVITY);
        else LIT(P, ANN_LAST_CHECK);
        if (fin[k]->kind == 1) P.in(fin[k]->raw);
        else P.aux(fin[k]->aux_off, fin[k]->aux_len);
      }
      LIT(P, RBRACE);
    }
    LIT(P, RBRACE);
  }
  if (!had_meta) {
    P.err = E_PARSE;
    return;
  }
  LIT(P, RBRACE);
}

struct GenCull {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;  // slot 1 is per-tuple, not one shared arena
  static const uint32_t CAP = 32;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { plan_cull(P, c); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    K k{c};
    uint32_t st = is_object(c.d[0], root(c.d[0])) ? E_OK : E_PARSE;
    return plan_objects(c, s, NOBJ, st, k);
  }
};

}  // namespace nbrm
