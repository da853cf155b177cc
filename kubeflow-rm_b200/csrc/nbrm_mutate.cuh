// nbrm_mutate.cuh — the ODH Notebook mutating webhook, minus its client reads.
//
// Reference: components/odh-notebook-controller/controllers/notebook_webhook.go
//   NotebookWebhook.Handle :234-302, InjectReconciliationLock :63-72,
//   InjectOAuthProxy :76-231, InjectCertConfig :421-535 (called from
//   CheckAndMountCACertBundle :373-419 once its ConfigMaps are found),
//   maybeRestartRunningNotebook :314-370; OAuthInjectionIsEnabled /
//   ServiceMeshIsEnabled notebook_controller.go:86,97.
// Golden: notebook_controller_test.go:602-658 with :1002-1086.
//
// GenMutate (slot 0 = the Notebook in the AdmissionRequest): lock on CREATE,
// CA bundle config when MUT_CA_BUNDLE, OAuth sidecar when the annotation asks
// for it; E_DENIED when OAuth and service mesh are both on (:273-274).
// GenRestartGuard (slot 0 = GenMutate's output, slot 1 = the request's object,
// slot 2 = its oldObject): the UPDATE-only rule that a running notebook's pod
// template is not changed by the webhook alone (:314-370).
#pragma once
#include "nbrm_cmp.cuh"
#include "nbrm_odh.cuh"

namespace nbrm {

enum MutFlags : uint32_t {
  MUT_CREATE = 1u,     // req.Operation == Create (:248)
  MUT_CA_BUNDLE = 2u,  // CheckAndMountCACertBundle found/created its ConfigMaps (:381-417)
};

NB_HD void plan_sidecar(Planner& P, const NbView& v, const EnvLits& env) {  // :78-165
  LIT(P, SIDECAR_HEAD);
  P.pool(env.oauth_image.off, env.oauth_image.len);
  LIT(P, SIDECAR_ARGS_1);
  P.in(inner(v.name));
  LIT(P, SIDECAR_ARGS_2);
  P.in(inner(v.name));
  LIT(P, SIDECAR_ARGS_3);
  if (v.a_logout.ok() && v.a_logout.len > 2) {  // :162-165
    LIT(P, SIDECAR_LOGOUT);
    P.in(inner(v.a_logout));
    LIT(P, QUOTE);
  }
  LIT(P, SIDECAR_TAIL);
}

template <uint32_t N>
NB_HD bool elem_named(const Doc& d, Span e, const char (&name)[N]) {
  return str_eq(d, FIND(d, e, "name"), name);
}

NB_HD bool elem_named_span(const Doc& d, Span e, Span name) {  // name: string value with quotes
  const Span n = FIND(d, e, "name");
  if (!is_string(d, name)) return !is_string(d, n) || n.len == 2;
  return is_string(d, n) && str_equal(d, n, d, name);
}

// The notebook image container with the CA bundle env vars and mount (:470-533).
NB_HD void plan_cert_container(Planner& P, const Doc& d, Span ctr) {
  bool have[5] = {false, false, false, false, false};
  const Span env = FIND(d, ctr, "env");
  {
    ArrIter it(d, env);
    Span e;
    while (it.next(d, e)) {
      const Span n = FIND(d, e, "name");
      if (str_eq(d, n, "PIP_CERT")) have[0] = true;
      else if (str_eq(d, n, "REQUESTS_CA_BUNDLE")) have[1] = true;
      else if (str_eq(d, n, "SSL_CERT_FILE")) have[2] = true;
      else if (str_eq(d, n, "PIPELINES_SSL_SA_CERTS")) have[3] = true;
      else if (str_eq(d, n, "GIT_SSL_CAINFO")) have[4] = true;
    }
  }
  const bool add_env = !(have[0] && have[1] && have[2] && have[3] && have[4]);
  LIT(P, LBRACE);
  Sep s;
  bool had_mounts = false;
  ObjIter it(d, ctr);
  Member m;
  while (it.next(d, m)) {
    if (KEY_IS(d, m, "env") && add_env) continue;  // rewritten below
    if (KEY_IS(d, m, "volumeMounts")) {
      had_mounts = true;
      s.next(P);
      LIT(P, KEY_VOLUMEMOUNTS);
      LIT(P, LBRACKET);
      Sep ms;
      bool replaced = false;
      ArrIter mi(d, m.val);
      Span e;
      while (mi.next(d, e)) {
        if (!replaced && elem_named(d, e, "trusted-ca")) {  // :508-514
          replaced = true;
          ms.next(P);
          LIT(P, MOUNT_TRUSTED_CA);
        } else {
          ms.in(P, d, e);
        }
      }
      if (!replaced) {
        ms.next(P);
        LIT(P, MOUNT_TRUSTED_CA);
      }
      LIT(P, RBRACKET);
      continue;
    }
    s.in(P, d, m.all);
  }
  if (add_env) {  // :476-496; the reference appends in Go map order, here in a fixed order
    s.next(P);
    LIT(P, KEY_ENV);
    LIT(P, LBRACKET);
    Sep es;
    ArrIter ei(d, env);
    Span e;
    while (ei.next(d, e)) es.in(P, d, e);
    if (!have[0]) { es.next(P); LIT(P, ENV_PIP_CERT); }
    if (!have[1]) { es.next(P); LIT(P, ENV_REQUESTS_CA); }
    if (!have[2]) { es.next(P); LIT(P, ENV_SSL_CERT); }
    if (!have[3]) { es.next(P); LIT(P, ENV_PIPELINES); }
    if (!have[4]) { es.next(P); LIT(P, ENV_GIT_SSL); }
    LIT(P, RBRACKET);
  }
  if (!had_mounts) {
    s.next(P);
    LIT(P, KEY_VOLUMEMOUNTS);
    LIT(P, LBRACKET);
    LIT(P, MOUNT_TRUSTED_CA);
    LIT(P, RBRACKET);
  }
  LIT(P, RBRACE);
}

NB_HD void plan_vol_oauth(Planner& P, const NbView& v) {
  LIT(P, VOL_OAUTH_HEAD);
  P.in(inner(v.name));
  LIT(P, VOL_OAUTH_TAIL);
}
NB_HD void plan_vol_tls(Planner& P, const NbView& v) {
  LIT(P, VOL_TLS_HEAD);
  P.in(inner(v.name));
  LIT(P, VOL_TLS_TAIL);
}

// PodSpec with the injections applied.
NB_HD void plan_mutated_podspec(Planner& P, const Doc& d, const NbView& v, const EnvLits& env,
                                bool cert, bool oauth, ViewOut& vw) {
  LIT(P, LBRACE);
  Sep ps;
  bool had_containers = false, had_volumes = false, had_sa = false;
  ObjIter it(d, v.podspec);
  Member m;
  while (it.next(d, m)) {
    if (KEY_IS(d, m, "containers") && (cert || oauth)) {
      had_containers = true;
      ps.next(P);
      LIT(P, KEY_CONTAINERS_ARR);
      const uint32_t arr0 = P.total;  // the readers of this Notebook get the extents of the array and of its
      LIT(P, LBRACKET);               // first element (VF_PORTS / VF_SUBJECTS, nbrm_nbc.cuh view_notebook)
      const uint32_t el0 = P.total;
      bool first_el = true;
      Sep cs;
      bool cert_done = !cert, proxy_done = !oauth;
      ArrIter ci(d, m.val);
      Span e;
      while (ci.next(d, e)) {
        // the CA bundle goes in first (:265), then the proxy replaces a container
        // named oauth-proxy (:170-176); one container cannot be both here
        if (!proxy_done && elem_named(d, e, "oauth-proxy")) {
          proxy_done = true;
          cs.next(P);
          plan_sidecar(P, v, env);
        } else if (!cert_done && elem_named_span(d, e, v.name)) {
          cert_done = true;
          cs.next(P);
          plan_cert_container(P, d, e);
        } else {
          cs.in(P, d, e);
        }
        if (first_el) vw.set(VF_SUBJECTS, el0, P.total);
        first_el = false;
      }
      if (!proxy_done) {  // :177-179
        cs.next(P);
        plan_sidecar(P, v, env);
        if (first_el) vw.set(VF_SUBJECTS, el0, P.total);
      }
      LIT(P, RBRACKET);
      vw.set(VF_PORTS, arr0, P.total);
      continue;
    }
    if (KEY_IS(d, m, "volumes") && (cert || oauth)) {
      had_volumes = true;
      ps.next(P);
      LIT(P, KEY_VOLUMES);
      LIT(P, LBRACKET);
      Sep vs;
      bool ca_done = !cert, oc_done = !oauth, tls_done = !oauth;
      ArrIter vi(d, m.val);
      Span e;
      while (vi.next(d, e)) {
        if (!ca_done && elem_named(d, e, "trusted-ca")) {  // :458-464
          ca_done = true;
          vs.next(P);
          LIT(P, VOL_TRUSTED_CA);
        } else if (!oc_done && elem_named(d, e, "oauth-config")) {  // :194-200
          oc_done = true;
          vs.next(P);
          plan_vol_oauth(P, v);
        } else if (!tls_done && elem_named(d, e, "tls-certificates")) {  // :217-223
          tls_done = true;
          vs.next(P);
          plan_vol_tls(P, v);
        } else {
          vs.in(P, d, e);
        }
      }
      if (!ca_done) { vs.next(P); LIT(P, VOL_TRUSTED_CA); }
      if (!oc_done) { vs.next(P); plan_vol_oauth(P, v); }
      if (!tls_done) { vs.next(P); plan_vol_tls(P, v); }
      LIT(P, RBRACKET);
      continue;
    }
    if (KEY_IS(d, m, "serviceAccountName") && oauth) {  // :229
      had_sa = true;
      ps.next(P);
      LIT(P, KEY_SERVICEACCOUNTNAME);
      put_str(P, v.name);
      continue;
    }
    ps.in(P, d, m.all);
  }
  if (!had_containers && oauth) {
    ps.next(P);
    LIT(P, KEY_CONTAINERS_ARR);
    const uint32_t arr0 = P.total;
    LIT(P, LBRACKET);
    const uint32_t el0 = P.total;
    plan_sidecar(P, v, env);
    vw.set(VF_SUBJECTS, el0, P.total);
    LIT(P, RBRACKET);
    vw.set(VF_PORTS, arr0, P.total);
  }
  if (!had_volumes && (cert || oauth)) {
    ps.next(P);
    LIT(P, KEY_VOLUMES);
    LIT(P, LBRACKET);
    Sep vs;
    if (cert) { vs.next(P); LIT(P, VOL_TRUSTED_CA); }
    if (oauth) {
      vs.next(P);
      plan_vol_oauth(P, v);
      vs.next(P);
      plan_vol_tls(P, v);
    }
    LIT(P, RBRACKET);
  }
  if (!had_sa && oauth) {
    ps.next(P);
    LIT(P, KEY_SERVICEACCOUNTNAME);
    put_str(P, v.name);
  }
  LIT(P, RBRACE);
}

// metadata with the annotation edits: add_lock sets the reconciliation lock
// (:63-72); pending = 0 leaves update-pending alone, 1 removes it (:292), 2 sets it (:290).
NB_HD void plan_mutated_metadata(Planner& P, const Doc& d, Span meta, bool add_lock, int pending, ViewOut& vw) {
  LIT(P, LBRACE);
  Sep ms;
  bool had_ann = false;
  ObjIter it(d, meta);
  Member m;
  while (it.next(d, m)) {
    if (!KEY_IS(d, m, "annotations")) {
      ms.in(P, d, m.all);
      if (KEY_IS(d, m, "labels")) vw.set(VF_LABELS, P.total - m.val.len, P.total);
      continue;
    }
    had_ann = true;
    if (!is_object(d, m.val) && !add_lock && pending != 2) {
      ms.in(P, d, m.all);
      vw.set(VF_ANNOTATIONS, P.total - m.val.len, P.total);
      continue;
    }
    ms.next(P);
    LIT(P, KEY_ANNOTATIONS);
    const uint32_t ann0 = P.total;
    LIT(P, LBRACE);
    Sep as;
    ObjIter ai(d, m.val);
    Member am;
    while (ai.next(d, am)) {
      if (add_lock && KEY_IS(d, am, "kubeflow-resource-stopped")) continue;
      if (pending && KEY_IS(d, am, "notebooks.opendatahub.io/update-pending")) continue;
      as.in(P, d, am.all);
    }
    if (add_lock) { as.next(P); LIT(P, ANN_LOCK); }
    if (pending == 2) { as.next(P); LIT(P, ANN_UPDATE_PENDING); }
    LIT(P, RBRACE);
    vw.set(VF_ANNOTATIONS, ann0, P.total);
  }
  if (!had_ann && (add_lock || pending == 2)) {
    ms.next(P);
    LIT(P, KEY_ANNOTATIONS);
    const uint32_t ann0 = P.total;
    LIT(P, LBRACE);
    Sep as;
    if (add_lock) { as.next(P); LIT(P, ANN_LOCK); }
    if (pending == 2) { as.next(P); LIT(P, ANN_UPDATE_PENDING); }
    LIT(P, RBRACE);
    vw.set(VF_ANNOTATIONS, ann0, P.total);
  }
  LIT(P, RBRACE);
}

// Notebook := Notebook with metadata / spec.template.spec replaced.  podspec_doc
// and podspec give the replacement PodSpec as a span of another slot (k), or
// plan it in place when podspec_k < 0.
struct MutatePlan {
  bool add_lock;
  int pending;
  bool cert, oauth;
  int podspec_k;    // -1: build the mutated PodSpec from slot 0; else copy span from slot k
  Span podspec_src;
};

NB_HD void plan_notebook_with(Planner& P, const Doc& d, const NbView& v, const EnvLits& env,
                              const MutatePlan& mp, ViewOut& vw) {
  const bool touch_spec = mp.podspec_k >= 0 || mp.cert || mp.oauth;
  LIT(P, LBRACE);
  Sep s;
  bool had_meta = false;
  ObjIter it(d, v.root);
  Member m;
  while (it.next(d, m)) {
    if (KEY_IS(d, m, "metadata") && is_object(d, m.val)) {
      had_meta = true;
      s.next(P);
      LIT(P, KEY_METADATA);
      const uint32_t m0 = P.total;
      plan_mutated_metadata(P, d, m.val, mp.add_lock, mp.pending, vw);
      vw.set(VF_META, m0, P.total);
    } else if (KEY_IS(d, m, "spec") && is_object(d, m.val) && touch_spec) {
      s.next(P);
      LIT(P, KEY_SPEC);
      const uint32_t s0 = P.total;
      LIT(P, LBRACE);
      Sep ss;
      ObjIter si(d, m.val);
      Member sm;
      while (si.next(d, sm)) {
        if (!(KEY_IS(d, sm, "template") && is_object(d, sm.val))) {
          ss.in(P, d, sm.all);
          continue;
        }
        ss.next(P);
        LIT(P, KEY_TEMPLATE);
        LIT(P, LBRACE);
        Sep ts;
        ObjIter ti(d, sm.val);
        Member tm;
        while (ti.next(d, tm)) {
          if (!KEY_IS(d, tm, "spec")) {
            ts.in(P, d, tm.all);
            continue;
          }
          ts.next(P);
          LIT(P, KEY_SPEC);
          const uint32_t p0 = P.total;
          if (mp.podspec_k >= 0) P.in(mp.podspec_src, (uint32_t)mp.podspec_k);
          else plan_mutated_podspec(P, d, v, env, mp.cert, mp.oauth, vw);
          vw.set(VF_TPOD, p0, P.total);
        }
        LIT(P, RBRACE);
      }
      LIT(P, RBRACE);
      vw.set(VF_SPEC, s0, P.total);
    } else {
      s.in(P, d, m.all);
      if (KEY_IS(d, m, "spec")) vw.set(VF_SPEC, P.total - m.val.len, P.total);
    }
  }
  if (!had_meta && (mp.add_lock || mp.pending == 2)) {
    s.next(P);
    LIT(P, KEY_METADATA);
    const uint32_t m0 = P.total;
    plan_mutated_metadata(P, d, mk(0, 0), mp.add_lock, mp.pending, vw);
    vw.set(VF_META, m0, P.total);
  }
  LIT(P, RBRACE);
  vw.set(VF_ROOT, 0, P.total);
}

struct GenMutate {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;  // slot 1 is per-tuple, not one shared arena
  static const uint32_t CAP = 224;
  struct K {
    DocCtx& c;
    const NbView& v;
    ViewTable views;
    uint32_t view_j;
    NB_HD void operator()(Planner& P, uint32_t) {
      const Doc& d = c.d[0];
      ViewOut vw;  // what the generators that read the mutated Notebook next will not have to look for
      vw.t = views;
      vw.j = view_j;
      MutatePlan mp;
      mp.add_lock = (c.flags & MUT_CREATE) != 0;
      mp.pending = (c.flags & MUT_CREATE) ? 1 : 0;  // Create: NoPendingUpdates -> delete (:319-322,:292)
      mp.cert = (c.flags & MUT_CA_BUNDLE) != 0;
      mp.oauth = oauth_enabled(d, v);
      mp.podspec_k = -1;
      mp.podspec_src = mk(0, 0);
      if (mp.oauth && mesh_enabled(d, v)) {  // :273-274
        P.err = E_DENIED;
        return;
      }
      if ((mp.cert || mp.oauth) && !is_object(d, v.podspec)) {
        P.err = E_PARSE;
        return;
      }
      plan_notebook_with(P, d, v, *c.env, mp, vw);
    }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    NbView v = view_notebook(c.d[0]);
    K k{c, v, s.views, s.view_j};
    return plan_objects(c, s, NOBJ, v.err, k);
  }
};

// maybeRestartRunningNotebook for an UPDATE (:314-370) plus the update-pending
// annotation (:288-293).  result = 1 when the webhook's own pod-template
// changes were held back.
struct GenRestartGuard {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;  // slot 1 is per-tuple, not one shared arena
  static const uint32_t CAP = 96;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) {
      const Doc& mut = c.d[0];
      const Doc& upd = c.d[1];
      const Doc& old = c.d[2];
      NbView mv = view_notebook(mut);
      if (mv.err) {
        P.err = mv.err;
        return;
      }
      MutatePlan mp;
      mp.add_lock = false;
      mp.pending = 1;
      mp.cert = mp.oauth = false;
      mp.podspec_k = -1;
      mp.podspec_src = mk(0, 0);
      // stopped (:325) or restarting (:331) notebooks take every change
      if (!mv.a_stopped.ok() && !mv.a_restart.ok()) {
        const Span up = FIND(upd, FIND(upd, FIND(upd, root(upd), "spec"), "template"), "spec");
        const Span op = FIND(old, FIND(old, FIND(old, root(old), "spec"), "template"), "spec");
        bool deep = false;
        const bool changed_by_webhook =
            semantic_equal(old, op, upd, up, &deep) && !semantic_equal(old, op, mut, mv.podspec, &deep);  // :353-362
        if (deep) {  // nested deeper than the compare follows: no answer, not "differs"
          P.err = E_LIMIT;
          return;
        }
        if (changed_by_webhook) {
          mp.pending = 2;          // :290
          mp.podspec_k = 1;        // :368 keep the pod template the request carried
          mp.podspec_src = up;
          c.result = 1;
          if (!up.ok()) {
            P.err = E_PARSE;
            return;
          }
        }
      }
      ViewOut vw;  // the guard's output is not one of the resident Notebooks the generators read
      vw.t.p = nullptr;
      vw.t.n = 0;
      vw.j = 0;
      plan_notebook_with(P, mut, mv, *c.env, mp, vw);
    }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    K k{c};
    if (c.d[0].n == 0) {  // the mutation stage rejected this Notebook: say why again
      s.nseg[0] = 0;
      s.lens[0] = 0;
      NbView uv = view_notebook(c.d[1]);
      if (!uv.err && oauth_enabled(c.d[1], uv) && mesh_enabled(c.d[1], uv)) return E_DENIED;
      return E_PARSE;
    }
    return plan_objects(c, s, NOBJ, E_OK, k);
  }
};

}  // namespace nbrm
