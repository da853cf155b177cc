// nbrm_nbc.cuh — planners for the notebook-controller generators.
//
// Reference behaviour (components/notebook-controller/controllers/
// notebook_controller.go): generateStatefulSet :408-484, setPrefixEnvVar
// :392-406, generateService :486-513, generateVirtualService :519-619,
// ctrl.SetControllerReference at :141,:183,:628.
#pragma once
#include "nbrm_gen.cuh"

namespace nbrm {

enum GenFlags : uint32_t {
  GEN_STS = 1u,
  GEN_SVC = 2u,
  GEN_VS = 4u,
  GEN_OWNER_REF = 8u,      // add the controller ownerReference (Reconcile :141,:183,:628)
  GEN_OWNER_V1 = 16u,      // ownerReference apiVersion kubeflow.org/v1 (ODH) instead of v1beta1
};

// Spans of a Notebook every planner needs.
struct NbView {
  Span root, meta, name, ns, uid, labels, annotations, podspec, containers, c0;
  // annotation values every generator asks for, found in ONE pass over the map
  Span a_stopped;   // kubeflow-resource-stopped (present when .ok())
  Span a_rewrite;   // notebooks.kubeflow.org/http-rewrite-uri
  Span a_headers;   // notebooks.kubeflow.org/http-headers-request-set
  Span a_oauth;     // notebooks.opendatahub.io/inject-oauth
  Span a_mesh;      // opendatahub.io/service-mesh
  Span a_logout;    // notebooks.opendatahub.io/oauth-logout-url
  Span a_restart;   // notebooks.opendatahub.io/notebook-restart
  bool l_statefulset, l_notebook_name, l_workbench;  // labels that override a default
  uint32_t err;
};

// One descent through the Notebook: every byte is read once (plus the values
// that are walked again by the planners that copy them).
NB_HD NbView view_notebook(const Doc& d) {
  NbView v;
  v.err = E_OK;
  v.root = v.meta = v.name = v.ns = v.uid = v.labels = v.annotations = mk(0, 0);
  v.podspec = v.containers = v.c0 = mk(0, 0);
  v.a_stopped = v.a_rewrite = v.a_headers = v.a_oauth = v.a_mesh = v.a_logout = v.a_restart = mk(0, 0);
  v.l_statefulset = v.l_notebook_name = v.l_workbench = false;
  Cursor c;
  c.i = skip_ws(d, 0);
  const uint32_t rs = c.i;
  if (!obj_enter(d, c)) {
    v.err = E_PARSE;
    return v;
  }
  Member m;
  while (obj_key(d, c, m)) {
    if (KEY_IS(d, m, "metadata")) {
      const uint32_t s0 = c.i;
      if (!obj_enter(d, c)) {
        val_skip(d, c);
        continue;
      }
      Member mm;
      while (obj_key(d, c, mm)) {
        if (KEY_IS(d, mm, "name")) {
          v.name = val_skip(d, c);
        } else if (KEY_IS(d, mm, "namespace")) {
          v.ns = val_skip(d, c);
        } else if (KEY_IS(d, mm, "uid")) {
          v.uid = val_skip(d, c);
        } else if (KEY_IS(d, mm, "labels")) {
          const uint32_t l0 = c.i;
          if (!obj_enter(d, c)) {
            val_skip(d, c);
            continue;
          }
          Member lm;
          while (obj_key(d, c, lm)) {
            if (KEY_IS(d, lm, "statefulset")) v.l_statefulset = true;
            else if (KEY_IS(d, lm, "notebook-name")) v.l_notebook_name = true;
            else if (KEY_IS(d, lm, "opendatahub.io/workbenches")) v.l_workbench = true;
            val_skip(d, c);
          }
          v.labels = mk(l0, c.i - l0);
        } else if (KEY_IS(d, mm, "annotations")) {
          const uint32_t a0 = c.i;
          if (!obj_enter(d, c)) {
            val_skip(d, c);
            continue;
          }
          Member am;
          while (obj_key(d, c, am)) {
            const Span val = val_skip(d, c);
            if (KEY_IS(d, am, "kubeflow-resource-stopped")) v.a_stopped = val;
            else if (KEY_IS(d, am, "notebooks.kubeflow.org/http-rewrite-uri")) v.a_rewrite = val;
            else if (KEY_IS(d, am, "notebooks.kubeflow.org/http-headers-request-set")) v.a_headers = val;
            else if (KEY_IS(d, am, "notebooks.opendatahub.io/inject-oauth")) v.a_oauth = val;
            else if (KEY_IS(d, am, "opendatahub.io/service-mesh")) v.a_mesh = val;
            else if (KEY_IS(d, am, "notebooks.opendatahub.io/oauth-logout-url")) v.a_logout = val;
            else if (KEY_IS(d, am, "notebooks.opendatahub.io/notebook-restart")) v.a_restart = val;
          }
          v.annotations = mk(a0, c.i - a0);
        } else {
          val_skip(d, c);
        }
      }
      v.meta = mk(s0, c.i - s0);
    } else if (KEY_IS(d, m, "spec")) {
      if (!obj_enter(d, c)) {
        val_skip(d, c);
        continue;
      }
      Member sm;
      while (obj_key(d, c, sm)) {
        if (!KEY_IS(d, sm, "template") || !obj_enter(d, c)) {
          val_skip(d, c);
          continue;
        }
        Member tm;
        while (obj_key(d, c, tm)) {
          if (!KEY_IS(d, tm, "spec")) {
            val_skip(d, c);
            continue;
          }
          const uint32_t p0 = c.i;
          if (!obj_enter(d, c)) {
            val_skip(d, c);
            continue;
          }
          Member pm;
          while (obj_key(d, c, pm)) {
            if (!KEY_IS(d, pm, "containers")) {
              val_skip(d, c);
              continue;
            }
            const uint32_t k0 = c.i;
            if (!arr_enter(d, c)) {
              val_skip(d, c);
              continue;
            }
            bool first = true;
            while (arr_elem(d, c)) {
              const Span e = val_skip(d, c);
              if (first && is_object(d, e)) v.c0 = e;
              first = false;
            }
            v.containers = mk(k0, c.i - k0);
          }
          v.podspec = mk(p0, c.i - p0);
        }
      }
    } else {
      val_skip(d, c);
    }
  }
  v.root = mk(rs, c.i - rs);
  if (!is_string(d, v.name)) v.name = mk(0, 0);
  if (!is_string(d, v.ns) || v.ns.len == 2) v.ns = mk(0, 0);
  if (!is_string(d, v.uid)) v.uid = mk(0, 0);
  return v;
}

// The same view of a Notebook another generator of this pass wrote (the webhook's mutated
// Notebook): it left the extents of metadata, labels, annotations, the pod spec, its containers
// and the first of them (VF_PORTS / VF_SUBJECTS stand for those two here), so only metadata's own
// members and the two small maps are walked.  Any of the large parts missing: the full walk.
NB_HD NbView view_notebook(const Doc& d, const ViewIn& vin) {
  if (!vin.ok()) return view_notebook(d);
  NbView v;
  v.root = vin.get(VF_ROOT);
  v.meta = vin.get(VF_META);
  v.podspec = vin.get(VF_TPOD);
  v.containers = vin.get(VF_PORTS);
  v.c0 = vin.get(VF_SUBJECTS);
  if (!v.root.ok() || !v.meta.ok() || !v.podspec.ok() || !v.containers.ok() || !v.c0.ok() || v.root.end() > d.n)
    return view_notebook(d);
  v.err = E_OK;
  v.labels = vin.get(VF_LABELS);
  v.annotations = vin.get(VF_ANNOTATIONS);
  v.name = v.ns = v.uid = mk(0, 0);
  v.a_stopped = v.a_rewrite = v.a_headers = v.a_oauth = v.a_mesh = v.a_logout = v.a_restart = mk(0, 0);
  v.l_statefulset = v.l_notebook_name = v.l_workbench = false;
  {
    ObjIter it(d, v.meta);
    Member m;
    while (it.next(d, m, v.labels, v.annotations)) {
      if (KEY_IS(d, m, "name")) v.name = m.val;
      else if (KEY_IS(d, m, "namespace")) v.ns = m.val;
      else if (KEY_IS(d, m, "uid")) v.uid = m.val;
    }
  }
  if (!is_object(d, v.labels)) v.labels = mk(0, 0);
  if (!is_object(d, v.annotations)) v.annotations = mk(0, 0);
  {
    ObjIter it(d, v.labels);
    Member m;
    while (it.next(d, m)) {
      if (KEY_IS(d, m, "statefulset")) v.l_statefulset = true;
      else if (KEY_IS(d, m, "notebook-name")) v.l_notebook_name = true;
      else if (KEY_IS(d, m, "opendatahub.io/workbenches")) v.l_workbench = true;
    }
  }
  {
    ObjIter it(d, v.annotations);
    Member m;
    while (it.next(d, m)) {
      if (KEY_IS(d, m, "kubeflow-resource-stopped")) v.a_stopped = m.val;
      else if (KEY_IS(d, m, "notebooks.kubeflow.org/http-rewrite-uri")) v.a_rewrite = m.val;
      else if (KEY_IS(d, m, "notebooks.kubeflow.org/http-headers-request-set")) v.a_headers = m.val;
      else if (KEY_IS(d, m, "notebooks.opendatahub.io/inject-oauth")) v.a_oauth = m.val;
      else if (KEY_IS(d, m, "opendatahub.io/service-mesh")) v.a_mesh = m.val;
      else if (KEY_IS(d, m, "notebooks.opendatahub.io/oauth-logout-url")) v.a_logout = m.val;
      else if (KEY_IS(d, m, "notebooks.opendatahub.io/notebook-restart")) v.a_restart = m.val;
    }
  }
  if (!is_object(d, v.c0)) v.c0 = mk(0, 0);
  if (!is_string(d, v.name)) v.name = mk(0, 0);
  if (!is_string(d, v.ns) || v.ns.len == 2) v.ns = mk(0, 0);
  if (!is_string(d, v.uid)) v.uid = mk(0, 0);
  return v;
}

NB_HD void put_str(Planner& P, Span s) {
  if (s.ok()) P.in(s);
  else LIT(P, EMPTY_STR);
}

// "name":<name>[,"namespace":<ns>][,"ownerReferences":[...]]  (after the head literal)
NB_HD void plan_meta_tail(Planner& P, const NbView& v, uint32_t flags) {
  if (v.ns.ok()) {
    LIT(P, NAMESPACE_KEY);
    P.in(v.ns);
  }
  if (flags & GEN_OWNER_REF) {
    if (flags & GEN_OWNER_V1) LIT(P, OWNER_V1);
    else LIT(P, OWNER_V1BETA1);
    put_str(P, v.name);
    LIT(P, OWNER_UID);
    put_str(P, v.uid);
    LIT(P, OWNER_TAIL);
  }
}

// strings.Contains(k, "kubectl") || strings.Contains(k, "notebook") (:449) in one
// pass over the key: the last eight bytes ride in a 64-bit register.
NB_HD bool key_has_kubectl_or_notebook(const Doc& d, Span k) {
  const unsigned long long NOTEBOOK = 0x6e6f7465626f6f6bull;  // "notebook"
  const unsigned long long KUBECTL = 0x006b75626563746cull;   // "kubectl"
  unsigned long long roll = 0;
  for (uint32_t j = 0; j < k.len; ++j) {
    roll = (roll << 8) | (unsigned char)d.at(k.off + j);
    if (j >= 7 && roll == NOTEBOOK) return true;
    if (j >= 6 && (roll & 0x00ffffffffffffffull) == KUBECTL) return true;
  }
  return false;
}

NB_HD bool env_has_nb_prefix(const Doc& d, Span env) {
  ArrIter it(d, env);
  Span e;
  while (it.next(d, e)) {
    if (str_eq(d, FIND(d, e, "name"), "NB_PREFIX")) return true;
  }
  return false;
}

NB_HD void plan_nb_prefix_elem(Planner& P, const NbView& v) {
  LIT(P, NB_PREFIX_OPEN);
  P.in(inner(v.ns));
  LIT(P, SLASH);
  P.in(inner(v.name));
  LIT(P, NB_PREFIX_CLOSE);
}

// generateStatefulSet (+ SetControllerReference when asked).
NB_HD void plan_statefulset(Planner& P, const Doc& d, const NbView& v, const EnvLits& env,
                            uint32_t flags, ViewOut& vw) {
  vw.clear();
  LIT(P, STS_HEAD);
  vw.set(VF_KIND, LIT_AFTER(STS_HEAD, "\"kind\":"), LIT_AFTER(STS_HEAD, "\"kind\":\"StatefulSet\""));
  put_str(P, v.name);
  plan_meta_tail(P, v, flags);
  vw.begin(VF_SPEC, P.total + LIT_AFTER(STS_SPEC_REPLICAS, "\"spec\":"));
  LIT(P, STS_SPEC_REPLICAS);
  vw.set(VF_REPLICAS, P.total, P.total + 1);
  if (v.a_stopped.ok()) LIT(P, ZERO);  // :410-412 metav1.HasAnnotation: presence only
  else LIT(P, ONE);
  vw.begin(VF_SELECTOR, P.total + LIT_AFTER(STS_SELECTOR, "\"selector\":"));
  LIT(P, STS_SELECTOR);
  put_str(P, v.name);
  vw.end(VF_SELECTOR, P.total + 2);
  vw.begin(VF_TLABELS, P.total + LIT_LEN(STS_TEMPLATE_LABELS) - 1);
  LIT(P, STS_TEMPLATE_LABELS);
  {  // :428-432 defaults, :441-444 Notebook labels win
    Sep s;
    if (!v.l_statefulset) {
      s.next(P);
      LIT(P, LBL_STATEFULSET);
      put_str(P, v.name);
    }
    if (!v.l_notebook_name) {
      s.next(P);
      LIT(P, LBL_NOTEBOOK_NAME);
      put_str(P, v.name);
    }
    if (!v.l_workbench) {
      s.next(P);
      LIT(P, LBL_WORKBENCH);
    }
    ObjIter it(d, v.labels);
    Member m;
    while (it.next(d, m)) s.in(P, d, m.all);
  }
  {  // :447-452 annotations whose key has neither "kubectl" nor "notebook"
    Sep s;
    ObjIter it(d, v.annotations);
    Member m;
    while (it.next(d, m)) {
      if (key_has_kubectl_or_notebook(d, inner(m.key))) continue;
      if (s.first) {
        vw.end(VF_TLABELS, P.total + 1);
        LIT(P, STS_ANNOTATIONS);
      }
      s.in(P, d, m.all);
    }
    if (s.first) vw.end(VF_TLABELS, P.total + 1);  // no annotations: the labels close in the next fragment
  }
  vw.begin(VF_TPOD, P.total + LIT_AFTER(STS_TEMPLATE_SPEC, "\"spec\":"));
  LIT(P, STS_TEMPLATE_SPEC);
  {  // PodSpec members, :435 deep copy with :454-482 edits
    Sep ps;
    bool has_sc = false;
    ObjIter it(d, v.podspec);
    Member m;
    while (it.next(d, m, v.containers)) {  // the containers' extent is known: not skipped through again
      if (KEY_IS(d, m, "containers")) {
        ps.next(P);
        LIT(P, KEY_CONTAINERS);
        Sep cs;
        bool has_wd = false, has_ports = false, has_env = false;
        ObjIter ci(d, v.c0);
        Member c;
        while (ci.next(d, c)) {
          if (KEY_IS(d, c, "workingDir")) {
            if (is_nil(d, c.val) || c.val.len == 2) continue;  // "" -> default, :456
            has_wd = true;
          } else if (KEY_IS(d, c, "ports")) {
            if (is_nil(d, c.val)) continue;  // nil -> default, :459
            has_ports = true;
          } else if (KEY_IS(d, c, "env")) {
            if (is_nil(d, c.val)) continue;
            has_env = true;
            if (is_array(d, c.val) && !env_has_nb_prefix(d, c.val)) {  // :402-405
              cs.next(P);
              LIT(P, KEY_ENV_OPEN);
              if (!is_empty_container(d, c.val)) {
                P.in(mk(c.val.off + 1, c.val.len - 2));
                LIT(P, COMMA);
              }
              plan_nb_prefix_elem(P, v);
              continue;
            }
          }
          cs.in(P, d, c.all);
        }
        if (!has_wd) {
          cs.next(P);
          LIT(P, KEY_WORKINGDIR_DEFAULT);
        }
        if (!has_ports) {
          cs.next(P);
          LIT(P, KEY_PORTS_DEFAULT);
        }
        if (!has_env) {
          cs.next(P);
          LIT(P, KEY_ENV_OPEN);
          plan_nb_prefix_elem(P, v);
        }
        LIT(P, RBRACE);
        // the other containers and the closing bracket, as they came
        P.in(mk(v.c0.end(), v.containers.end() - v.c0.end()));
        continue;
      }
      if (KEY_IS(d, m, "securityContext")) {
        if (is_nil(d, m.val)) continue;  // nil pointer, :476
        has_sc = true;
      }
      ps.in(P, d, m.all);
    }
    if (env.add_fsgroup && !has_sc) {  // :475-481
      ps.next(P);
      LIT(P, KEY_FSGROUP_DEFAULT);
    }
  }
  LIT(P, STS_TAIL);  // }}}} closes the PodSpec, the template, the spec, the object
  vw.end(VF_TPOD, P.total - 3);
  vw.end(VF_SPEC, P.total - 1);
  vw.set(VF_ROOT, 0, P.total);
}

// generateService.
NB_HD void plan_service(Planner& P, const Doc& d, const NbView& v, uint32_t flags, ViewOut& vw) {
  vw.clear();
  LIT(P, SVC_HEAD);
  vw.set(VF_KIND, LIT_AFTER(SVC_HEAD, "\"kind\":"), LIT_AFTER(SVC_HEAD, "\"kind\":\"Service\""));
  put_str(P, v.name);
  plan_meta_tail(P, v, flags);
  vw.begin(VF_SPEC, P.total + LIT_AFTER(SVC_PORTS, "\"spec\":"));
  vw.begin(VF_PORTS, P.total + LIT_AFTER(SVC_PORTS, "\"ports\":"));
  LIT(P, SVC_PORTS);
  P.in(inner(v.name));
  LIT(P, SVC_TARGETPORT);
  Span ports = FIND(d, v.c0, "ports");
  if (is_nil(d, ports)) {  // :490 nil test
    LIT(P, PORT_8888);
  } else {
    ArrIter it(d, ports);
    Span e;
    if (!it.next(d, e)) {
      P.err = E_PORTS_EMPTY;  // :491 would index an empty slice
      return;
    }
    Span cp = FIND(d, e, "containerPort");
    if (cp.ok() && !is_null(d, cp)) P.in(cp);
    else LIT(P, ZERO);
  }
  vw.end(VF_PORTS, P.total + 2);
  vw.begin(VF_SELECTOR, P.total + LIT_AFTER(SVC_SELECTOR, "\"selector\":"));
  LIT(P, SVC_SELECTOR);
  put_str(P, v.name);
  vw.end(VF_SELECTOR, P.total + 1);
  LIT(P, SVC_TAIL);
  vw.end(VF_SPEC, P.total - 1);
  vw.set(VF_ROOT, 0, P.total);
}

NB_HD int hexval(char c) {
  if (c >= '0' && c <= '9') return c - '0';
  if (c >= 'a' && c <= 'f') return c - 'a' + 10;
  if (c >= 'A' && c <= 'F') return c - 'A' + 10;
  return -1;
}

// Undo one level of JSON string escaping.  Returns bytes written, or
// 0xffffffff when the escapes are malformed or the slice is too small.
NB_HD uint32_t unescape(const Doc& d, Span s, char* out, uint32_t cap) {
  uint32_t o = 0, i = s.off, e = s.end();
  while (i < e) {
    if (o + 4 > cap) return 0xffffffffu;
    char c = d.at(i);
    if (c != '\\') {
      out[o++] = c;
      ++i;
      continue;
    }
    if (i + 1 >= e) return 0xffffffffu;
    char x = d.at(i + 1);
    i += 2;
    switch (x) {
      case '"': out[o++] = '"'; break;
      case '\\': out[o++] = '\\'; break;
      case '/': out[o++] = '/'; break;
      case 'b': out[o++] = '\b'; break;
      case 'f': out[o++] = '\f'; break;
      case 'n': out[o++] = '\n'; break;
      case 'r': out[o++] = '\r'; break;
      case 't': out[o++] = '\t'; break;
      case 'u': {
        if (i + 4 > e) return 0xffffffffu;
        uint32_t cp = 0;
        for (int k = 0; k < 4; ++k) {
          int h = hexval(d.at(i + k));
          if (h < 0) return 0xffffffffu;
          cp = cp * 16 + (uint32_t)h;
        }
        i += 4;
        if (cp >= 0xD800 && cp < 0xDC00 && i + 6 <= e && d.at(i) == '\\' && d.at(i + 1) == 'u') {
          uint32_t lo = 0;
          bool okh = true;
          for (int k = 0; k < 4; ++k) {
            int h = hexval(d.at(i + 2 + k));
            if (h < 0) okh = false;
            lo = lo * 16 + (uint32_t)(h < 0 ? 0 : h);
          }
          if (okh && lo >= 0xDC00 && lo < 0xE000) {
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
            i += 6;
          }
        }
        if (cp >= 0xD800 && cp < 0xE000) cp = 0xFFFD;  // lone surrogate
        if (cp < 0x80) {
          out[o++] = (char)cp;
        } else if (cp < 0x800) {
          out[o++] = (char)(0xC0 | (cp >> 6));
          out[o++] = (char)(0x80 | (cp & 0x3F));
        } else if (cp < 0x10000) {
          out[o++] = (char)(0xE0 | (cp >> 12));
          out[o++] = (char)(0x80 | ((cp >> 6) & 0x3F));
          out[o++] = (char)(0x80 | (cp & 0x3F));
        } else {
          out[o++] = (char)(0xF0 | (cp >> 18));
          out[o++] = (char)(0x80 | ((cp >> 12) & 0x3F));
          out[o++] = (char)(0x80 | ((cp >> 6) & 0x3F));
          out[o++] = (char)(0x80 | (cp & 0x3F));
        }
        break;
      }
      default:
        return 0xffffffffu;
    }
  }
  return o;
}

// End index of the strictly valid JSON string starting at i, or 0.
NB_HD uint32_t strict_string_end(const Doc& d, uint32_t i) {
  if (i >= d.n || d.at(i) != '"') return 0;
  uint32_t j = i + 1;
  while (j < d.n) {
    unsigned char c = (unsigned char)d.at(j);
    if (c == '"') return j + 1;
    if (c < 0x20) return 0;
    if (c == '\\') {
      if (j + 1 >= d.n) return 0;
      char x = d.at(j + 1);
      if (x == 'u') {
        if (j + 6 > d.n) return 0;
        for (int k = 2; k < 6; ++k)
          if (hexval(d.at(j + k)) < 0) return 0;
        j += 6;
      } else if (x == '"' || x == '\\' || x == '/' || x == 'b' || x == 'f' || x == 'n' ||
                 x == 'r' || x == 't') {
        j += 2;
      } else {
        return 0;
      }
    } else {
      ++j;
    }
  }
  return 0;
}

// headers.request.set from the http-headers-request-set annotation (:566-579):
// the annotation must decode to a flat object of strings (null counts as ""),
// anything else gives the empty map.  Emits the members (no braces).
NB_HD void plan_headers_set(Planner& P, const Doc& d, const NbView& v, Aux& aux) {
  Span a = v.a_headers;
  if (!is_string(d, a) || a.len == 2) return;
  char* out = aux.p + aux.used;
  uint32_t n = unescape(d, inner(a), out, aux.cap - aux.used);
  if (n == 0xffffffffu) return;
  Doc t;
  t.p = out;
  t.n = n;
  // pass 1: validate strictly; pass 2: emit
  for (int pass = 0; pass < 2; ++pass) {
    uint32_t i = skip_ws(t, 0);
    if (i >= t.n || t.at(i) != '{') return;
    i = skip_ws(t, i + 1);
    bool first = true;
    if (i < t.n && t.at(i) == '}') {
      ++i;
    } else {
      for (;;) {
        uint32_t ke = strict_string_end(t, i);
        if (!ke) return;
        uint32_t ks = i;
        i = skip_ws(t, ke);
        if (i >= t.n || t.at(i) != ':') return;
        i = skip_ws(t, i + 1);
        uint32_t vs = i, ve;
        bool isnull = false;
        if (i + 4 <= t.n && t.at(i) == 'n' && t.at(i + 1) == 'u' && t.at(i + 2) == 'l' &&
            t.at(i + 3) == 'l') {
          ve = i + 4;
          isnull = true;
        } else {
          ve = strict_string_end(t, i);
          if (!ve) return;
        }
        if (pass == 1) {
          if (!first) LIT(P, COMMA);
          P.aux(aux.abs + aux.used + ks, ke - ks);
          LIT(P, COLON);
          if (isnull) LIT(P, EMPTY_STR);
          else P.aux(aux.abs + aux.used + vs, ve - vs);
        }
        first = false;
        i = skip_ws(t, ve);
        if (i < t.n && t.at(i) == ',') {
          i = skip_ws(t, i + 1);
          continue;
        }
        if (i < t.n && t.at(i) == '}') {
          ++i;
          break;
        }
        return;
      }
    }
    i = skip_ws(t, i);
    if (i != t.n) return;
  }
  aux.used += n;
}

// generateVirtualService.
NB_HD void plan_virtualservice(Planner& P, const Doc& d, const NbView& v, const EnvLits& env,
                               uint32_t flags, Aux& aux, ViewOut& vw) {
  vw.clear();
  LIT(P, VS_HEAD);  // ... "name":"notebook-
  vw.set(VF_KIND, LIT_AFTER(VS_HEAD, "\"kind\":"), LIT_AFTER(VS_HEAD, "\"kind\":\"VirtualService\""));
  P.in(inner(v.ns));
  LIT(P, DASH);
  P.in(inner(v.name));
  LIT(P, QUOTE);
  plan_meta_tail(P, v, flags);
  vw.begin(VF_SPEC, P.total + LIT_AFTER(VS_GATEWAYS, "\"spec\":"));
  LIT(P, VS_GATEWAYS);
  P.pool(env.istio_gateway.off, env.istio_gateway.len);
  LIT(P, VS_HOSTS);
  P.pool(env.istio_host.off, env.istio_host.len);
  LIT(P, VS_HTTP_SET);
  plan_headers_set(P, d, v, aux);
  LIT(P, VS_MATCH);
  P.in(inner(v.ns));
  LIT(P, SLASH);
  P.in(inner(v.name));
  LIT(P, VS_REWRITE);
  Span rw = v.a_rewrite;
  if (is_string(d, rw) && rw.len > 2) {  // :533
    P.in(rw);
  } else {
    LIT(P, VS_REWRITE_DEFAULT);
    P.in(inner(v.ns));
    LIT(P, SLASH);
    P.in(inner(v.name));
    LIT(P, SLASH_QUOTE);
  }
  LIT(P, VS_ROUTE);
  P.in(inner(v.name));
  LIT(P, DOT);
  P.in(inner(v.ns));
  LIT(P, DOT_SVC_DOT);
  P.pool(env.cluster_domain.off, env.cluster_domain.len);
  LIT(P, VS_TAIL);
  vw.end(VF_SPEC, P.total - 1);
  vw.set(VF_ROOT, 0, P.total);
}

// The notebook-controller generator: StatefulSet, Service, VirtualService.
struct GenNbc {
  static const bool VIEWS = true;  // leaves a view of each object for the compare stage (nbrm_plan.cuh)
  static const uint32_t NOBJ = 3;
  static const bool SHARED1 = false;  // slot 1 is per-tuple, not one shared arena
  static const uint32_t CAP = 96;
  struct K {
    DocCtx& c;
    const NbView& v;
    ViewTable views;
    uint32_t view_j;
    NB_HD void operator()(Planner& P, uint32_t k) {
      if (!(c.flags & (1u << k))) return;
      ViewOut vw;
      vw.t = views;
      vw.j = view_j + k;
      if (k == 0) plan_statefulset(P, c.d[0], v, *c.env, c.flags, vw);
      else if (k == 1) plan_service(P, c.d[0], v, c.flags, vw);
      else plan_virtualservice(P, c.d[0], v, *c.env, c.flags, c.aux, vw);
    }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    NbView v = view_notebook(c.d[0], c.view[0]);
    uint32_t st = v.err;
    if (!st && (!is_object(c.d[0], v.podspec) || !v.c0.ok())) st = E_NO_CONTAINERS;
    K k{c, v, s.views, s.view_j};
    return plan_objects(c, s, NOBJ, st, k);
  }
};

}  // namespace nbrm
