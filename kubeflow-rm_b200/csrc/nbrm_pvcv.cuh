// nbrm_pvcv.cuh — what one PVCViewer reconcile writes.
//
// Reference: components/pvcviewer-controller/controllers/pvcviewer_controller.go — Reconcile
// :96-146 (commonLabels :121-125), reconcileDeployment :149-208, reconcileService :211-250,
// reconcileVirtualService :252-336, the Affinity generateAffinity ends with :445-464,
// ctrl.SetControllerReference :198,:240,:326.  Each reconcile* function Gets its object,
// overwrites the fields it owns and Creates (NotFound) or Updates what it then holds; the
// generator emits the body of that write.
// Input slot 0: PVCViewer CRs.  Slot 1 (shared arena, optional): the three Gets' results,
// document 3 i + k = found Deployment / Service / VirtualService of viewer i (empty: NotFound).
// Slot 2 (optional): per viewer the node its ReadWriteOnce PVC is mounted on, as a JSON string
// ("" or empty: none) — the PVC Get and the Pod List that find it (:372-443) stay with the caller;
// used only for rwoScheduling viewers whose Deployment is being created (:170).
// Output: Deployment, Service, VirtualService; the last two have length 0 when spec.networking
// is the zero value (:212,:253).  An object that has another controller gives E_CONFLICT
// (SetControllerReference's AlreadyOwnedError).
#pragma once
#include "nbrm_cmp.cuh"
#include "nbrm_nbc.cuh"

namespace nbrm {

struct PvView {
  Span name, ns, uid, podspec, net, tport, base, rewrite, timeout;
  bool rwo, net_zero;
};

NB_HD bool pv_empty_str(const Doc& d, Span s) { return is_nil(d, s) || (is_string(d, s) && s.len == 2); }

NB_HD PvView view_pvcviewer(const Doc& d) {
  PvView v;
  const Span rt = root(d);
  const Span meta = FIND(d, rt, "metadata");
  v.name = FIND(d, meta, "name");
  v.ns = FIND(d, meta, "namespace");
  v.uid = FIND(d, meta, "uid");
  const Span spec = FIND(d, rt, "spec");
  v.podspec = FIND(d, spec, "podSpec");
  v.net = FIND(d, spec, "networking");
  v.tport = FIND(d, v.net, "targetPort");
  v.base = FIND(d, v.net, "basePrefix");
  v.rewrite = FIND(d, v.net, "rewrite");
  v.timeout = FIND(d, v.net, "timeout");
  const Span rwo = FIND(d, spec, "rwoScheduling");
  v.rwo = rwo.len == 4 && d.at(rwo.off) == 't';
  if (!is_string(d, v.name)) v.name = mk(0, 0);
  if (!is_string(d, v.ns)) v.ns = mk(0, 0);
  if (!is_string(d, v.uid)) v.uid = mk(0, 0);
  // Networking{} == zero value: an IntOrString of 0 / "" and three empty strings
  const bool tp_zero = pv_empty_str(d, v.tport) || (v.tport.len == 1 && d.at(v.tport.off) == '0');
  v.net_zero = tp_zero && pv_empty_str(d, v.base) && pv_empty_str(d, v.rewrite) && pv_empty_str(d, v.timeout);
  return v;
}

NB_HD void plan_pv_labels(Planner& P, const PvView& v) {
  LIT(P, PV_LABELS_HEAD);
  put_str(P, v.name);
  LIT(P, PV_LABELS_INSTANCE);
  P.in(inner(v.name));
  LIT(P, PV_LABELS_TAIL);
}

NB_HD void plan_pv_owner(Planner& P, const PvView& v) {
  LIT(P, PV_OWNER);
  put_str(P, v.name);
  LIT(P, OWNER_UID);
  put_str(P, v.uid);
  LIT(P, PV_OWNER_TAIL);
}

// The found object k of this viewer (n == 0: NotFound).
NB_HD Doc pv_found(const DocCtx& c, uint32_t k, uint32_t& abs) {
  Doc f;
  f.p = nullptr;
  f.n = 0;
  abs = 0;
  const uint64_t j = (uint64_t)c.index * 3u + k;
  if (c.shared && j < c.shared_n) {
    abs = (uint32_t)c.shared_off[j];
    f.p = c.shared + c.shared_off[j];
    f.n = (uint32_t)(c.shared_off[j + 1] - c.shared_off[j]);
  }
  return f;
}

// metadata of the written object.  Create: name, namespace, labels, our ownerReference.  Update:
// the found metadata with ownerReferences upserted as controllerutil.SetControllerReference does.
NB_HD void plan_pv_meta(Planner& P, const Doc& d, const PvView& v, const Doc& f, Span fmeta, bool ns_always) {
  LIT(P, PV_META_KEY);
  if (!f.n) {
    LIT(P, PV_META_NAME);
    P.in(inner(v.name));
    LIT(P, QUOTE);
    if (v.ns.ok() && (ns_always || v.ns.len > 2)) {
      LIT(P, NAMESPACE_KEY);
      P.in(v.ns);
    } else if (ns_always) {
      LIT(P, NAMESPACE_KEY);
      LIT(P, EMPTY_STR);
    }
    LIT(P, PV_LABELS_KEY);
    plan_pv_labels(P, v);
    LIT(P, PV_OWNERS_KEY);
    plan_pv_owner(P, v);
    LIT(P, PV_OWNERS_TAIL);
    return;
  }
  LIT(P, LBRACE);
  Sep s;
  {
    ObjIter it(f, fmeta);
    Member m;
    while (it.next(f, m)) {
      if (str_eq(f, m.key, "ownerReferences")) continue;
      s.next(P);
      P.in(m.all, 1);
    }
  }
  s.next(P);
  LIT(P, PV_OWNERS_KEY_BARE);
  bool placed = false;
  Sep rs;
  ArrIter it(f, FIND(f, fmeta, "ownerReferences"));
  Span r;
  while (it.next(f, r)) {
    const Span ctl = FIND(f, r, "controller");
    const Span uid = FIND(f, r, "uid");
    const bool same_uid = is_string(f, uid) && v.uid.ok() ? str_equal(f, uid, d, v.uid) : (!v.uid.ok() && pv_empty_str(f, uid));
    if (ctl.len == 4 && f.at(ctl.off) == 't' && !same_uid) {
      P.err = E_CONFLICT;
      return;
    }
    // the same owner: group kubeflow.org, kind PVCViewer, this name
    const Span av = FIND(f, r, "apiVersion"), kd = FIND(f, r, "kind"), nm = FIND(f, r, "name");
    bool group = false;
    if (is_string(f, av)) {
      const Span in = inner(av);
      group = bytes_eq(f, in, "kubeflow.org") ||
              (in.len > 13 && bytes_eq(f, mk(in.off, 13), "kubeflow.org/"));
    }
    const bool same_name = is_string(f, nm) && v.name.ok() ? str_equal(f, nm, d, v.name) : (!v.name.ok() && pv_empty_str(f, nm));
    rs.next(P);
    if (!placed && group && str_eq(f, kd, "PVCViewer") && same_name) {
      plan_pv_owner(P, v);
      placed = true;
    } else {
      P.in(r, 1);
    }
  }
  if (!placed) {
    rs.next(P);
    plan_pv_owner(P, v);
  }
  LIT(P, PV_OWNERS_TAIL);
}

// Root members of a found object other than metadata (and spec): copied as they are.
template <bool SKIP_SPEC>
NB_HD void plan_pv_rest(Planner& P, const Doc& f) {
  ObjIter it(f, root(f));
  Member m;
  while (it.next(f, m)) {
    if (str_eq(f, m.key, "metadata") || (SKIP_SPEC && str_eq(f, m.key, "spec"))) continue;
    P.in(m.all, 1);
    LIT(P, COMMA);
  }
}

template <uint32_t N0, uint32_t N1, uint32_t N2>
NB_HD void plan_pv_spec_rest(Planner& P, const Doc& f, Span fspec, const char (&a)[N0], const char (&b)[N1],
                             const char (&c)[N2]) {
  ObjIter it(f, fspec);
  Member m;
  while (it.next(f, m)) {
    if (str_eq(f, m.key, a) || str_eq(f, m.key, b) || str_eq(f, m.key, c)) continue;
    P.in(m.all, 1);
    LIT(P, COMMA);
  }
}

NB_HD void plan_pv_deployment(Planner& P, const DocCtx& c, const PvView& v) {
  const Doc& d = c.d[0];
  uint32_t abs;
  const Doc f = pv_found(c, 0, abs);
  P.base[1] = abs;
  if (f.n && !is_object(f, root(f))) {
    P.err = E_PARSE;
    return;
  }
  LIT(P, LBRACE);
  if (f.n) plan_pv_rest<true>(P, f);
  else LIT(P, PV_DEP_TYPE);
  const Span fmeta = f.n ? FIND(f, root(f), "metadata") : mk(0, 0);
  plan_pv_meta(P, d, v, f, fmeta, false);
  if (P.err) return;
  LIT(P, PV_SPEC_KEY);
  const Span fspec = f.n ? FIND(f, root(f), "spec") : mk(0, 0);
  if (f.n) plan_pv_spec_rest(P, f, fspec, "selector", "template", "strategy");
  LIT(P, PV_DEP_SELECTOR);
  plan_pv_labels(P, v);
  LIT(P, PV_DEP_TEMPLATE);
  plan_pv_labels(P, v);
  LIT(P, PV_DEP_PODSPEC);
  {  // viewer.Spec.PodSpec with the affinity of :168-196
    Sep s;
    ObjIter it(d, v.podspec);
    Member m;
    while (it.next(d, m)) {
      if (str_eq(d, m.key, "affinity")) continue;
      s.in(P, d, m.all);
    }
    Span faff = mk(0, 0);
    if (f.n) faff = FIND(f, FIND(f, FIND(f, fspec, "template"), "spec"), "affinity");
    const Span node = root(c.d[2]);
    if (v.rwo && !f.n && is_string(c.d[2], node) && node.len > 2) {
      s.next(P);
      LIT(P, PV_AFFINITY_KEY);
      LIT(P, TB_AFF_HEAD);
      P.in(node, 2);
      LIT(P, TB_AFF_TAIL);
    } else if (f.n && !is_nil(f, faff)) {
      s.next(P);
      LIT(P, PV_AFFINITY_KEY);
      P.in(faff, 1);
    }
  }
  LIT(P, PV_DEP_TAIL);
}

NB_HD void plan_pv_service(Planner& P, const DocCtx& c, const PvView& v) {
  if (v.net_zero) return;
  const Doc& d = c.d[0];
  uint32_t abs;
  const Doc f = pv_found(c, 1, abs);
  P.base[1] = abs;
  if (f.n && !is_object(f, root(f))) {
    P.err = E_PARSE;
    return;
  }
  LIT(P, LBRACE);
  if (f.n) plan_pv_rest<true>(P, f);
  else LIT(P, PV_SVC_TYPE);
  plan_pv_meta(P, d, v, f, f.n ? FIND(f, root(f), "metadata") : mk(0, 0), false);
  if (P.err) return;
  LIT(P, PV_SPEC_KEY);
  if (f.n) plan_pv_spec_rest(P, f, FIND(f, root(f), "spec"), "type", "selector", "ports");
  LIT(P, PV_SVC_SELECTOR);
  plan_pv_labels(P, v);
  LIT(P, PV_SVC_PORTS);
  if (is_nil(d, v.tport)) LIT(P, ZERO);
  else P.in(v.tport);
  LIT(P, PV_SVC_TAIL);
}

NB_HD void plan_pv_virtualservice(Planner& P, const DocCtx& c, const PvView& v) {
  if (v.net_zero) return;
  const Doc& d = c.d[0];
  uint32_t abs;
  const Doc f = pv_found(c, 2, abs);
  P.base[1] = abs;
  if (f.n && !is_object(f, root(f))) {
    P.err = E_PARSE;
    return;
  }
  LIT(P, LBRACE);
  if (f.n) plan_pv_rest<true>(P, f);
  else LIT(P, PV_VS_TYPE);
  plan_pv_meta(P, d, v, f, f.n ? FIND(f, root(f), "metadata") : mk(0, 0), true);
  if (P.err) return;
  LIT(P, PV_VS_SPEC);
  P.pool(c.env->istio_gateway.off, c.env->istio_gateway.len);
  LIT(P, PV_VS_MATCH);
  const bool has_base = is_string(d, v.base);
  if (has_base) P.in(inner(v.base));
  LIT(P, SLASH);
  P.in(inner(v.ns));
  LIT(P, SLASH);
  P.in(inner(v.name));
  LIT(P, PV_VS_REWRITE);
  if (!pv_empty_str(d, v.rewrite) && is_string(d, v.rewrite)) {
    P.in(v.rewrite);
  } else {
    LIT(P, QUOTE);
    if (has_base) P.in(inner(v.base));
    LIT(P, SLASH);
    P.in(inner(v.ns));
    LIT(P, SLASH);
    P.in(inner(v.name));
    LIT(P, SLASH_QUOTE);
  }
  LIT(P, PV_VS_ROUTE);
  P.in(inner(v.name));
  LIT(P, DOT);
  P.in(inner(v.ns));
  LIT(P, PV_VS_TIMEOUT);
  if (!pv_empty_str(d, v.timeout) && is_string(d, v.timeout)) P.in(v.timeout);
  else LIT(P, NULL_LIT);
  LIT(P, PV_VS_TAIL);
}

struct GenPvcViewer {
  static const uint32_t NOBJ = 3;
  static const bool SHARED1 = true;
  static const uint32_t CAP = 160;
  struct K {
    DocCtx& c;
    const PvView& v;
    NB_HD void operator()(Planner& P, uint32_t k) {
      if (k == 0) plan_pv_deployment(P, c, v);
      else if (k == 1) plan_pv_service(P, c, v);
      else plan_pv_virtualservice(P, c, v);
    }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    const Doc& d = c.d[0];
    const uint32_t i0 = skip_ws(d, 0);
    const uint32_t st = (i0 < d.n && d.at(i0) == '{') ? E_OK : E_PARSE;
    PvView v = view_pvcviewer(d);
    K k{c, v};
    return plan_objects(c, s, NOBJ, st, k);
  }
};

// ---- reconcileStatus :339-369 -------------------------------------------------------------
// Slot 0: the viewer, slot 1: its Deployment (empty: the Get failed).  Output: the status
// r.Client.Status().Update sends: url from the networking, ready = replicas == readyReplicas,
// the Deployment's first condition appended when the last one differs.
NB_HD void plan_pv_status(Planner& P, DocCtx& c) {
  const Doc& d = c.d[0];
  const Doc& dep = c.d[1];
  const PvView v = view_pvcviewer(d);
  const Span conds = FIND(d, FIND(d, root(d), "status"), "conditions");
  Span last = mk(0, 0);
  uint32_t n_old = 0;
  {
    ArrIter it(d, conds);
    Span e;
    while (it.next(d, e)) {
      last = e;
      ++n_old;
    }
  }
  Span d0 = mk(0, 0);
  bool ready = false;
  if (dep.n) {
    const Span dr = root(dep);
    const Span rep = FIND(dep, FIND(dep, dr, "spec"), "replicas");
    if (is_nil(dep, rep)) {  // *deployment.Spec.Replicas :357
      P.err = E_NIL_REPLICAS;
      return;
    }
    const Span dst = FIND(dep, dr, "status");
    const Span rr = FIND(dep, dst, "readyReplicas");
    ready = is_nil(dep, rr) ? (rep.len == 1 && dep.at(rep.off) == '0') : spans_eq(dep, rep, dep, rr);
    ArrIter it(dep, FIND(dep, dst, "conditions"));
    Span e;
    if (it.next(dep, e)) d0 = e;
  }
  bool deep = false;
  const bool append = d0.ok() && (n_old == 0 || !json_equal(d, last, dep, d0, &deep));  // :360-363
  if (deep) {
    P.err = E_LIMIT;
    return;
  }
  LIT(P, LBRACE);
  if (n_old || append) {
    LIT(P, PVS_CONDS);
    Sep s;
    ArrIter it(d, conds);
    Span e;
    while (it.next(d, e)) s.in(P, d, e);
    if (append) {
      s.next(P);
      P.in(d0, 1);
    }
    LIT(P, PVS_CONDS_CLOSE);
  }
  if (ready) LIT(P, PVS_READY_TRUE);
  else LIT(P, PVS_READY_FALSE);
  if (!v.net_zero) {
    LIT(P, PVS_URL);
    if (is_string(d, v.base)) P.in(inner(v.base));
    LIT(P, SLASH);
    P.in(inner(v.ns));
    LIT(P, SLASH);
    P.in(inner(v.name));
    LIT(P, SLASH_QUOTE);
  }
  LIT(P, RBRACE);
}

struct GenPvStatus {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;
  static const uint32_t CAP = 48;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { plan_pv_status(P, c); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    const Doc& d = c.d[0];
    const uint32_t i0 = skip_ws(d, 0);
    const uint32_t st = (i0 < d.n && d.at(i0) == '{') ? E_OK : E_PARSE;
    K k{c};
    return plan_objects(c, s, NOBJ, st, k);
  }
};

// ---- the viewer's webhook ------------------------------------------------------------------
// Reference: pvcviewer-controller/api/v1alpha1/pvcviewer_webhook.go — Default :70-146 (the
// defaults file of DEFAULT_POD_SPEC_PATH :78-91 arrives through the environment, split by the
// host into its members other than `volumes` and its volumes), validate :153-178.
// Output: the viewer as Default leaves it.  result: 0 = valid, else the message of validate:
// 1 "PVC name must be specified", 2 "PodSpec must be specified", 3 "PVC <pvc> must be used in
// the podSpec".
enum PvInvalid : uint32_t { PV_VALID = 0, PV_NO_PVC = 1, PV_NO_PODSPEC = 2, PV_PVC_UNUSED = 3 };

// reflect.DeepEqual(podSpec, corev1.PodSpec{}) on the decoded struct: absent, null, {} or only nulls
NB_HD bool pv_podspec_zero(const Doc& d, Span ps) {
  if (is_nil(d, ps)) return true;
  if (!is_object(d, ps)) return false;
  ObjIter it(d, ps);
  Member m;
  while (it.next(d, m))
    if (!is_null(d, m.val)) return false;
  return true;
}

NB_HD void plan_pv_default(Planner& P, DocCtx& c) {
  const Doc& d = c.d[0];
  const Span rt = root(d);
  const Span spec = FIND(d, rt, "spec");
  const Span ps = FIND(d, spec, "podSpec");
  const Span pvc = FIND(d, spec, "pvc");
  const bool pvc_ok = is_string(d, pvc) && pvc.len > 2;
  const bool zero = pv_podspec_zero(d, ps);
  c.result = PV_VALID;
  if (!pvc_ok) {
    c.result = PV_NO_PVC;
  } else if (!zero) {  // the PVC must be one of the pod's volumes (:165-176)
    bool used = false;
    ArrIter it(d, FIND(d, ps, "volumes"));
    Span e;
    while (it.next(d, e)) {
      const Span cn = FIND(d, FIND(d, e, "persistentVolumeClaim"), "claimName");
      if (is_string(d, cn) && str_equal(d, cn, d, pvc)) used = true;
    }
    if (!used) c.result = PV_PVC_UNUSED;
  }
  if (!zero) {
    P.in(rt);
    return;
  }
  // the viewer with spec.podSpec filled in
  const PvView v = view_pvcviewer(d);
  LIT(P, LBRACE);
  {
    ObjIter it(d, rt);
    Member m;
    while (it.next(d, m)) {
      if (str_eq(d, m.key, "spec")) continue;
      P.in(m.all);
      LIT(P, COMMA);
    }
  }
  LIT(P, PV_WH_SPEC);
  {
    ObjIter it(d, spec);
    Member m;
    while (it.next(d, m)) {
      if (str_eq(d, m.key, "podSpec")) continue;
      P.in(m.all);
      LIT(P, COMMA);
    }
  }
  LIT(P, PV_WH_PODSPEC);
  if (c.env->pv_default_rest.len || c.env->pv_default_volumes.len || c.env->pv_default_set) {
    P.pool(c.env->pv_default_rest.off, c.env->pv_default_rest.len);
    if (c.env->pv_default_rest.len) LIT(P, COMMA);
    LIT(P, PV_WH_VOLUMES);
    P.pool(c.env->pv_default_volumes.off, c.env->pv_default_volumes.len);
    if (c.env->pv_default_volumes.len) LIT(P, COMMA);
  } else {
    LIT(P, PV_WH_CONTAINER);
    if (is_string(d, v.base)) P.in(inner(v.base));
    LIT(P, SLASH);
    P.in(inner(v.ns));
    LIT(P, SLASH);
    P.in(inner(v.name));
    LIT(P, PV_WH_CONTAINER_TAIL);
  }
  LIT(P, PV_WH_VOLUME);
  if (is_string(d, pvc)) P.in(pvc);
  else LIT(P, EMPTY_STR);
  LIT(P, PV_WH_TAIL);
}

struct GenPvcViewerDefault {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;
  static const uint32_t CAP = 96;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { plan_pv_default(P, c); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    const Doc& d = c.d[0];
    const uint32_t i0 = skip_ws(d, 0);
    const uint32_t st = (i0 < d.n && d.at(i0) == '{') ? E_OK : E_PARSE;
    K k{c};
    return plan_objects(c, s, NOBJ, st, k);
  }
};

}  // namespace nbrm
