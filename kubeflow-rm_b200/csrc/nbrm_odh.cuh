// nbrm_odh.cuh — planners for the objects OpenshiftNotebookReconciler builds.
//
// Reference (components/odh-notebook-controller/controllers/):
//   NewNotebookNetworkPolicy notebook_network.go:131-173, NewOAuthNetworkPolicy
//   :176-210, NewNotebookRoute notebook_route.go:34-62, NewNotebookOAuthRoute
//   notebook_oauth.go:253-259, NewNotebookServiceAccount :49-64,
//   NewNotebookOAuthService :116-140, NewNotebookOAuthSecret :190-210,
//   NewRoleBinding notebook_rbac.go:36-58; which of them a reconcile builds:
//   notebook_controller.go:188-232; OAuthInjectionIsEnabled :86,
//   ServiceMeshIsEnabled :97.
#pragma once
#include "nbrm_nbc.cuh"

namespace nbrm {

enum OdhFlags : uint32_t {
  ODH_RBAC = 32u,  // SET_PIPELINE_RBAC == "true" and the Role exists (notebook_rbac.go:61,144)
};

static const uint32_t ODH_OUT_PER_DOC = 7;  // ctrl-np, oauth-np, route, sa, service, secret, rolebinding

// strconv.ParseBool on an annotation value (raw JSON string span): true for
// 1 t T TRUE true True, anything else false (errors are dropped, :88,:99).
NB_HD bool annotation_true(const Doc& d, Span v) {
  if (!is_string(d, v)) return false;
  Span s = inner(v);
  return bytes_eq(d, s, "true") || bytes_eq(d, s, "1") || bytes_eq(d, s, "t") ||
         bytes_eq(d, s, "T") || bytes_eq(d, s, "TRUE") || bytes_eq(d, s, "True");
}

NB_HD bool oauth_enabled(const Doc& d, const NbView& v) {
  return annotation_true(d, v.a_oauth);
}
NB_HD bool mesh_enabled(const Doc& d, const NbView& v) {
  return annotation_true(d, v.a_mesh);
}

NB_HD void plan_labels_notebook_name(Planner& P, const NbView& v, ViewOut& vw) {
  vw.begin(VF_LABELS, P.total + LIT_AFTER(LABELS_NOTEBOOK_NAME, "\"labels\":"));
  LIT(P, LABELS_NOTEBOOK_NAME);
  put_str(P, v.name);
  vw.end(VF_LABELS, P.total + 1);  // the next fragment starts with the closing brace
}

NB_HD void plan_network_policy(Planner& P, const NbView& v, const EnvLits& env, uint32_t flags,
                               bool oauth_np, ViewOut& vw) {
  LIT(P, NP_HEAD);
  vw.set(VF_KIND, LIT_AFTER(NP_HEAD, "\"kind\":"), LIT_AFTER(NP_HEAD, "\"kind\":\"NetworkPolicy\""));
  P.in(inner(v.name));
  if (oauth_np) LIT(P, NP_OAUTH_SUFFIX);
  else LIT(P, NP_CTRL_SUFFIX);
  plan_meta_tail(P, v, flags);
  vw.begin(VF_SPEC, P.total + LIT_AFTER(NP_SPEC, "\"spec\":"));
  LIT(P, NP_SPEC);
  put_str(P, v.name);
  if (oauth_np) {
    LIT(P, NP_OAUTH_TAIL);
  } else {
    LIT(P, NP_CTRL_RULE);
    P.pool(env.ctrl_namespace.off, env.ctrl_namespace.len);
    LIT(P, NP_CTRL_TAIL);
  }
  vw.end(VF_SPEC, P.total - 1);
  vw.set(VF_ROOT, 0, P.total);
}

NB_HD void plan_route(Planner& P, const NbView& v, uint32_t flags, bool oauth_route, ViewOut& vw) {
  LIT(P, ROUTE_HEAD);
  vw.set(VF_KIND, LIT_AFTER(ROUTE_HEAD, "\"kind\":"), LIT_AFTER(ROUTE_HEAD, "\"kind\":\"Route\""));
  put_str(P, v.name);
  plan_meta_tail(P, v, flags);
  plan_labels_notebook_name(P, v, vw);
  vw.begin(VF_SPEC, P.total + LIT_AFTER(ROUTE_SPEC_TO, "\"spec\":"));
  LIT(P, ROUTE_SPEC_TO);
  P.in(inner(v.name));
  if (oauth_route) {
    vw.end(VF_SPEC, P.total + LIT_AFTER(ROUTE_OAUTH_TAIL, "\"wildcardPolicy\":\"None\"}"));
    LIT(P, ROUTE_OAUTH_TAIL);
  } else {
    LIT(P, ROUTE_PLAIN_PORT);
    P.in(inner(v.name));
    vw.end(VF_SPEC, P.total + LIT_AFTER(ROUTE_PLAIN_TAIL, "\"wildcardPolicy\":\"None\"}"));
    LIT(P, ROUTE_PLAIN_TAIL);
  }
  vw.set(VF_ROOT, 0, P.total);
}

NB_HD void plan_service_account(Planner& P, const NbView& v, uint32_t flags, ViewOut& vw) {
  LIT(P, SA_HEAD);
  vw.set(VF_KIND, LIT_AFTER(SA_HEAD, "\"kind\":"), LIT_AFTER(SA_HEAD, "\"kind\":\"ServiceAccount\""));
  put_str(P, v.name);
  plan_meta_tail(P, v, flags);
  plan_labels_notebook_name(P, v, vw);
  LIT(P, SA_ANNOTATION);
  P.in(inner(v.name));
  LIT(P, SA_TAIL);
  vw.set(VF_ROOT, 0, P.total);
}

NB_HD void plan_oauth_service(Planner& P, const NbView& v, uint32_t flags, ViewOut& vw) {
  LIT(P, TLSSVC_HEAD);
  vw.set(VF_KIND, LIT_AFTER(TLSSVC_HEAD, "\"kind\":"), LIT_AFTER(TLSSVC_HEAD, "\"kind\":\"Service\""));
  P.in(inner(v.name));
  LIT(P, TLS_SUFFIX);
  plan_meta_tail(P, v, flags);
  plan_labels_notebook_name(P, v, vw);
  LIT(P, TLSSVC_ANNOTATION);
  P.in(inner(v.name));
  LIT(P, TLSSVC_SPEC);
  put_str(P, v.name);
  LIT(P, RBRACE3);
  vw.set(VF_ROOT, 0, P.total);
}

NB_HD char b64c(uint32_t x) {
  return (char)(x < 26 ? 'A' + x : x < 52 ? 'a' + (x - 26) : x < 62 ? '0' + (x - 52) : x == 62 ? '+' : '/');
}

// base64 of n bytes (n a multiple of 3, or 16 with "==" padding) into out.
NB_HD uint32_t base64(const unsigned char* in, uint32_t n, char* out) {
  uint32_t o = 0, i = 0;
  for (; i + 3 <= n; i += 3) {
    uint32_t w = (in[i] << 16) | (in[i + 1] << 8) | in[i + 2];
    out[o++] = b64c(w >> 18);
    out[o++] = b64c((w >> 12) & 63);
    out[o++] = b64c((w >> 6) & 63);
    out[o++] = b64c(w & 63);
  }
  if (n - i == 1) {
    uint32_t w = in[i] << 16;
    out[o++] = b64c(w >> 18);
    out[o++] = b64c((w >> 12) & 63);
    out[o++] = '=';
    out[o++] = '=';
  } else if (n - i == 2) {
    uint32_t w = (in[i] << 16) | (in[i + 1] << 8);
    out[o++] = b64c(w >> 18);
    out[o++] = b64c((w >> 12) & 63);
    out[o++] = b64c((w >> 6) & 63);
    out[o++] = '=';
  }
  return o;
}

// cookie_secret = base64(base64(16 seed bytes)), 32 characters (:192-195).  The
// seed comes from the caller's CSPRNG; nothing random is drawn on the device.
NB_HD void plan_oauth_secret(Planner& P, const NbView& v, uint32_t flags, const unsigned char* seed16,
                             Aux& aux, ViewOut& vw) {
  LIT(P, SECRET_HEAD);
  vw.set(VF_KIND, LIT_AFTER(SECRET_HEAD, "\"kind\":"), LIT_AFTER(SECRET_HEAD, "\"kind\":\"Secret\""));
  P.in(inner(v.name));
  LIT(P, SECRET_SUFFIX);
  plan_meta_tail(P, v, flags);
  plan_labels_notebook_name(P, v, vw);
  LIT(P, SECRET_DATA);
  if (aux.used + 64 <= aux.cap) {
    char tmp[24];
    unsigned char s[16];
    for (int k = 0; k < 16; ++k) s[k] = seed16 ? seed16[k] : 0;
    base64(s, 16, tmp);
    char* out = aux.p + aux.used;
    uint32_t n = base64((const unsigned char*)tmp, 24, out);
    P.aux(aux.abs + aux.used, n);
    aux.used += n;
  } else {
    P.err = E_LIMIT;
  }
  LIT(P, SECRET_TAIL);
  vw.set(VF_ROOT, 0, P.total);
}

NB_HD void plan_rolebinding(Planner& P, const NbView& v, uint32_t flags, ViewOut& vw) {
  LIT(P, RB_HEAD);
  vw.set(VF_KIND, LIT_AFTER(RB_HEAD, "\"kind\":"), LIT_AFTER(RB_HEAD, "\"kind\":\"RoleBinding\""));
  P.in(inner(v.name));
  LIT(P, QUOTE);
  plan_meta_tail(P, v, flags);
  plan_labels_notebook_name(P, v, vw);
  vw.begin(VF_SUBJECTS, P.total + LIT_AFTER(RB_SUBJECTS, "\"subjects\":"));
  LIT(P, RB_SUBJECTS);
  put_str(P, v.name);
  if (v.ns.ok()) {
    LIT(P, NAMESPACE_KEY);
    P.in(v.ns);
  }
  vw.end(VF_SUBJECTS, P.total + 2);  // "}]" opens the last fragment
  LIT(P, RB_TAIL);
  vw.set(VF_ROOT, 0, P.total);
}

// One ODH reconcile's objects; slot k of ODH_OUT_PER_DOC, or nothing when the
// reference would not build that object for this Notebook.
NB_HD void plan_odh_object(Planner& P, const Doc& d, const NbView& v, const EnvLits& env,
                           uint32_t flags, uint32_t k, bool oauth, bool mesh,
                           const unsigned char* seed16, Aux& aux, ViewOut& vw) {
  vw.clear();
  switch (k) {
    case 0: plan_network_policy(P, v, env, flags, false, vw); break;
    case 1: if (!mesh) plan_network_policy(P, v, env, flags, true, vw); break;
    case 2: if (!mesh) plan_route(P, v, flags, oauth, vw); break;
    case 3: if (!mesh && oauth) plan_service_account(P, v, flags, vw); break;
    case 4: if (!mesh && oauth) plan_oauth_service(P, v, flags, vw); break;
    case 5: if (!mesh && oauth) plan_oauth_secret(P, v, flags, seed16, aux, vw); break;
    case 6: if (flags & ODH_RBAC) plan_rolebinding(P, v, flags, vw); break;
  }
}

// The ODH reconcile generator.  extra = 16 seed bytes per Notebook (or null).
struct GenOdh {
  static const bool VIEWS = true;  // leaves a view of each object for the write decision (nbrm_plan.cuh)
  static const uint32_t NOBJ = ODH_OUT_PER_DOC;
  static const bool SHARED1 = false;  // slot 1 is per-tuple, not one shared arena
  static const uint32_t CAP = 32;
  struct K {
    DocCtx& c;
    const NbView& v;
    bool oauth, mesh;
    ViewTable views;
    uint32_t view_j;
    NB_HD void operator()(Planner& P, uint32_t k) {
      const unsigned char* seed =
          c.extra ? (const unsigned char*)c.extra + (size_t)c.index * 16 : nullptr;
      ViewOut vw;
      vw.t = views;
      vw.j = view_j + k;
      plan_odh_object(P, c.d[0], v, *c.env, c.flags, k, oauth, mesh, seed, c.aux, vw);
    }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    NbView v = view_notebook(c.d[0], c.view[0]);
    K k{c, v, false, false, s.views, s.view_j};
    if (!v.err) {
      k.oauth = oauth_enabled(c.d[0], v);
      k.mesh = mesh_enabled(c.d[0], v);
    }
    return plan_objects(c, s, NOBJ, v.err, k);
  }
};

}  // namespace nbrm
