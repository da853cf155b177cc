// nbrm_tb.cuh — the objects one Tensorboard reconcile builds.
//
// Reference: components/tensorboard-controller/controllers/tensorboard_controller.go —
// generateDeployment :167-299 (log path kinds :380-412, the node affinity generateNodeAffinity
// leaves behind :445-465, rwoPVCScheduling :471-482), generateService :301-319,
// generateVirtualService :321-378, ctrl.SetControllerReference :94,:103,:112.
// Input slot 0: Tensorboard CRs.  Input slot 1 (optional): per Tensorboard the name of the node
// a running Pod mounts its ReadWriteOnce PVC on, as a JSON string ("" or an empty document:
// none) — the PVC Get and the Pod List that find it are I/O and stay with the caller.
// Output: Deployment, Service, VirtualService.  A required environment variable that is not set
// (TENSORBOARD_IMAGE, ISTIO_GATEWAY, ISTIO_HOST: getEnvVariable :535-541) or an invalid
// RWO_PVC_SCHEDULING on a PVC path gives E_ENV and no objects: the reference's Reconcile returns
// the error (having written the Deployment and the Service first when it is the VirtualService's
// variables that are missing; the batch form fails the whole tuple).
#pragma once
#include "nbrm_cmp.cuh"
#include "nbrm_nbc.cuh"

namespace nbrm {

struct TbView {
  Span name, ns, uid, labels, logs;
};

NB_HD TbView view_tensorboard(const Doc& d) {
  TbView v;
  const Span rt = root(d);
  const Span meta = FIND(d, rt, "metadata");
  v.name = FIND(d, meta, "name");
  v.ns = FIND(d, meta, "namespace");
  v.uid = FIND(d, meta, "uid");
  v.labels = FIND(d, meta, "labels");
  v.logs = FIND(d, FIND(d, rt, "spec"), "logspath");
  if (!is_string(d, v.name)) v.name = mk(0, 0);
  if (!is_string(d, v.ns) || v.ns.len == 2) v.ns = mk(0, 0);
  if (!is_string(d, v.uid)) v.uid = mk(0, 0);
  if (!is_string(d, v.logs)) v.logs = mk(0, 0);
  return v;
}

// The log path, decoded into the document's aux slice, and what kind it is.
struct TbPath {
  const char* b;  // decoded bytes
  uint32_t n;
  bool raw;       // no escapes: decoded offsets are offsets into the input string too
  bool cloud, gcs, pvc;
  uint32_t name0, name1;  // pvc://<name>/<sub>: decoded byte ranges
  uint32_t sub0, sub1;
};

template <uint32_t N>
NB_HD bool tb_prefix(const TbPath& p, const char (&lit)[N]) {
  if (p.n < N - 1) return false;
  for (uint32_t k = 0; k < N - 1; ++k)
    if (p.b[k] != lit[k]) return false;
  return true;
}

NB_HD bool tb_path(const Doc& d, Span logs, Aux& aux, TbPath& p) {
  const Span in = inner(logs);
  p.raw = true;
  for (uint32_t k = 0; k < in.len; ++k)
    if (d.at(in.off + k) == '\\') p.raw = false;
  if (aux.used + in.len + 4 > aux.cap) return false;
  char* out = aux.p + aux.used;
  const uint32_t n = in.len ? unescape(d, in, out, aux.cap - aux.used) : 0u;
  if (n == 0xffffffffu) return false;
  aux.used += n;
  p.b = out;
  p.n = n;
  p.gcs = tb_prefix(p, "gs://");                                               // :384-386
  p.cloud = p.gcs || tb_prefix(p, "s3://") || tb_prefix(p, "/cns/");            // :380-382
  p.pvc = tb_prefix(p, "pvc://");                                              // :388-390
  p.name0 = p.name1 = p.sub0 = p.sub1 = 0;
  if (p.pvc) {  // extractPVCName :392-400, extractPVCSubPath :402-410
    p.name0 = 6;
    uint32_t e = 6;
    while (e < n && out[e] != '/') ++e;
    p.name1 = e;
    if (e < n && e + 1 < n) {
      p.sub0 = e + 1;
      p.sub1 = n;
    }
  }
  return true;
}

// Decoded bytes [a, b) of the path as the inside of a JSON string: straight from the input when
// the path had no escapes, else escaped again into aux.
NB_HD void tb_put_piece(Planner& P, const Doc& d, Span logs, const TbPath& p, uint32_t a, uint32_t b, Aux& aux) {
  if (b <= a) return;
  if (p.raw) {
    P.in(mk(logs.off + 1 + a, b - a));
    return;
  }
  const uint32_t at = aux.used;
  for (uint32_t k = a; k < b; ++k) {
    const unsigned char c = (unsigned char)p.b[k];
    if (aux.used + 6 > aux.cap) {
      P.err = E_LIMIT;
      return;
    }
    char* o = aux.p + aux.used;
    if (c == '"' || c == '\\') {
      o[0] = '\\';
      o[1] = (char)c;
      aux.used += 2;
    } else if (c < 0x20) {
      const char hex[] = "0123456789abcdef";
      o[0] = '\\';
      o[1] = 'u';
      o[2] = '0';
      o[3] = '0';
      o[4] = hex[c >> 4];
      o[5] = hex[c & 15];
      aux.used += 6;
    } else {
      o[0] = (char)c;
      aux.used += 1;
    }
  }
  P.aux(aux.abs + at, aux.used - at);
}

// "name":<name>[,"namespace":<ns>][,"ownerReferences":[...]] after a head literal that ends in "name":
NB_HD void plan_tb_meta(Planner& P, const TbView& v, uint32_t flags) {
  put_str(P, v.name);
  if (v.ns.ok()) {
    LIT(P, NAMESPACE_KEY);
    P.in(v.ns);
  }
  if (flags & GEN_OWNER_REF) {
    LIT(P, TB_OWNER);
    put_str(P, v.name);
    LIT(P, OWNER_UID);
    put_str(P, v.uid);
    LIT(P, OWNER_TAIL);
  }
}

enum TbEnvBits : uint32_t { TB_ENV_IMAGE = 1u, TB_ENV_GATEWAY = 2u, TB_ENV_HOST = 4u, TB_ENV_RWO = 8u, TB_ENV_RWO_BAD = 16u };

NB_HD void plan_tb_deployment(Planner& P, const Doc& d, const Doc& dn, const TbView& v, const EnvLits& env,
                              uint32_t flags, Aux& aux) {
  TbPath p;
  if (!tb_path(d, v.logs, aux, p)) {
    P.err = E_LIMIT;
    return;
  }
  const bool pvc_volume = !p.cloud;  // :178: a PVC holds the logs (pvc://name/sub, or the legacy tb-volume)
  if (pvc_volume && (env.tb_env & TB_ENV_RWO_BAD)) {  // rwoPVCScheduling is only asked on this path (:207)
    P.err = E_ENV;
    return;
  }
  LIT(P, TB_DEP_HEAD);
  plan_tb_meta(P, v, flags);
  LIT(P, TB_DEP_SPEC);
  put_str(P, v.name);
  LIT(P, TB_DEP_TLABELS);
  put_str(P, v.name);
  {  // :252-256 the CR's labels, "app" overwritten
    ObjIter it(d, v.labels);
    Member m;
    while (it.next(d, m)) {
      if (str_eq(d, m.key, "app")) continue;
      LIT(P, COMMA);
      P.in(m.all);
    }
  }
  LIT(P, TB_DEP_AFFINITY);
  bool with_node = false;
  if (pvc_volume && (env.tb_env & TB_ENV_RWO)) {  // :207-232
    const Span node = root(dn);
    if (is_string(dn, node) && node.len > 2) {
      LIT(P, TB_AFF_HEAD);
      P.in(node, 1);
      LIT(P, TB_AFF_TAIL);
      with_node = true;
    }
  }
  if (!with_node) {
    LIT(P, LBRACE);
    LIT(P, RBRACE);
  }
  LIT(P, TB_DEP_CONTAINER);
  if (env.tb_image.len > 2) {  // omitempty
    LIT(P, TB_IMAGE_KEY);
    P.pool(env.tb_image.off, env.tb_image.len);
  }
  LIT(P, TB_DEP_CMD);
  if (pvc_volume && p.pvc) LIT(P, TB_LOGDIR_PVC);  // mountpath = /tensorboard_logs/ (:183)
  else tb_put_piece(P, d, v.logs, p, 0, p.n, aux);
  LIT(P, TB_DEP_ARGS_TAIL);
  if (pvc_volume) {
    LIT(P, TB_MOUNT_PVC);
    if (p.pvc) LIT(P, TB_LOGDIR_PVC);
    else tb_put_piece(P, d, v.logs, p, 0, p.n, aux);
    if (p.pvc && p.sub1 > p.sub0) {
      LIT(P, TB_MOUNT_SUBPATH);
      tb_put_piece(P, d, v.logs, p, p.sub0, p.sub1, aux);
    }
    LIT(P, TB_MOUNT_PVC_TAIL);
    if (p.pvc) tb_put_piece(P, d, v.logs, p, p.name0, p.name1, aux);
    else LIT(P, TB_VOLUME_LEGACY);
    LIT(P, TB_VOLUME_PVC_TAIL);
  } else if (p.gcs) {
    LIT(P, TB_GCS);
  } else {
    LIT(P, TB_NO_VOLUMES);
  }
}

NB_HD void plan_tb_service(Planner& P, const TbView& v, uint32_t flags) {
  LIT(P, TB_SVC_HEAD);
  plan_tb_meta(P, v, flags);
  LIT(P, TB_SVC_SPEC);
  put_str(P, v.name);
  LIT(P, TB_SVC_PORTS);
  P.in(inner(v.name));
  LIT(P, TB_SVC_TAIL);
}

NB_HD void plan_tb_virtualservice(Planner& P, const TbView& v, const EnvLits& env, uint32_t flags) {
  LIT(P, TB_VS_HEAD);
  plan_tb_meta(P, v, flags);
  LIT(P, TB_VS_HOSTS);
  P.pool(env.tb_host.off, env.tb_host.len);
  LIT(P, TB_VS_GATEWAYS);
  P.pool(env.tb_gateway.off, env.tb_gateway.len);
  LIT(P, TB_VS_MATCH);
  P.in(inner(v.ns));
  LIT(P, SLASH);
  P.in(inner(v.name));
  LIT(P, TB_VS_ROUTE);
  P.in(inner(v.name));
  LIT(P, DOT);
  P.in(inner(v.ns));
  LIT(P, TB_VS_TAIL);
}

struct GenTensorboard {
  static const uint32_t NOBJ = 3;
  static const bool SHARED1 = false;
  static const uint32_t CAP = 64;
  struct K {
    DocCtx& c;
    const TbView& v;
    NB_HD void operator()(Planner& P, uint32_t k) {
      if (k == 0) plan_tb_deployment(P, c.d[0], c.d[1], v, *c.env, c.flags, c.aux);
      else if (k == 1) plan_tb_service(P, v, c.flags);
      else plan_tb_virtualservice(P, v, *c.env, c.flags);
    }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    const Doc& d = c.d[0];
    const uint32_t i0 = skip_ws(d, 0);
    uint32_t st = (i0 < d.n && d.at(i0) == '{') ? E_OK : E_PARSE;
    const uint32_t need = TB_ENV_IMAGE | TB_ENV_GATEWAY | TB_ENV_HOST;
    if (!st && (c.env->tb_env & need) != need) st = E_ENV;
    TbView v = view_tensorboard(d);
    K k{c, v};
    return plan_objects(c, s, NOBJ, st, k);
  }
};

// ---- after the Get of the Deployment -----------------------------------------------------
// Reference: Deployment :484-510 with CopyDeploymentSetFields :512-533 (labels of the found
// object that the desired one does not carry with the same value, replicas, affinity), and
// the status update of Reconcile :119-152.  Slot 0: Tensorboard CR, slot 1: the desired
// Deployment (GenTensorboard's), slot 2: the found one (empty: NotFound — the caller Creates
// the desired object and writes no status, :491-498,:124-129).  Output 0: the body of the
// Update (the found object with labels, replicas and affinity of the desired one), length 0
// when CopyDeploymentSetFields says nothing changed.  Output 1: the Tensorboard's status as
// r.Status().Update sends it (:148).  result: 1 = the Deployment is updated.
NB_HD void plan_tb_deployment_update(Planner& P, DocCtx& c) {
  const Doc& from = c.d[1];
  const Doc& to = c.d[2];
  if (!to.n) return;
  const Span fr = root(from), tr = root(to);
  if (!is_object(from, fr) || !is_object(to, tr)) {
    P.err = E_PARSE;
    return;
  }
  const Span flabels = FIND(from, FIND(from, fr, "metadata"), "labels");
  const Span tmeta = FIND(to, tr, "metadata");
  const Span fspec = FIND(from, fr, "spec"), tspec = FIND(to, tr, "spec");
  const Span frep = FIND(from, fspec, "replicas"), trep = FIND(to, tspec, "replicas");
  const Span faff = FIND(from, FIND(from, FIND(from, fspec, "template"), "spec"), "affinity");
  const Span ttmpl = FIND(to, tspec, "template");
  const Span tpod = FIND(to, ttmpl, "spec");
  const Span taff = FIND(to, tpod, "affinity");
  if (is_nil(from, frep) || is_nil(to, trep)) {  // *from.Spec.Replicas != *to.Spec.Replicas :521
    P.err = E_NIL_REPLICAS;
    return;
  }
  bool deep = false;
  bool update = map_field_differs(from, flabels, to, FIND(to, tmeta, "labels"));  // :514-518
  if (!spans_eq(from, frep, to, trep)) update = true;
  const bool fnil = is_nil(from, faff), tnil = is_nil(to, taff);
  if (fnil != tnil || (!fnil && !json_equal(from, faff, to, taff, &deep))) update = true;  // :526-529
  if (deep) {
    P.err = E_LIMIT;
    return;
  }
  if (!update) return;
  c.result |= 1u;
  // the found object, with the three fields of the desired one
  LIT(P, LBRACE);
  Sep s;
  ObjIter it(to, tr);
  Member m;
  bool had_meta = false, had_spec = false;
  while (it.next(to, m)) {
    if (str_eq(to, m.key, "metadata") && is_object(to, m.val)) {
      had_meta = true;
      s.next(P);
      LIT(P, TBU_META);
      Sep ms;
      ObjIter mi(to, m.val);
      Member mm;
      while (mi.next(to, mm)) {
        if (str_eq(to, mm.key, "labels")) continue;
        ms.in(P, to, mm.all, 2);
      }
      if (!is_nil(from, flabels)) {  // to.Labels = from.Labels :519
        ms.next(P);
        LIT(P, TBU_LABELS);
        P.in(flabels, 1);
      }
      LIT(P, RBRACE);
    } else if (str_eq(to, m.key, "spec") && is_object(to, m.val)) {
      had_spec = true;
      s.next(P);
      LIT(P, PV_WH_SPEC);
      Sep ss;
      ObjIter si(to, m.val);
      Member sm;
      bool had_tmpl = false;
      while (si.next(to, sm)) {
        if (str_eq(to, sm.key, "replicas")) continue;
        if (str_eq(to, sm.key, "template") && is_object(to, sm.val)) {
          had_tmpl = true;
          ss.next(P);
          LIT(P, TBU_TEMPLATE);
          Sep ts;
          ObjIter ti(to, sm.val);
          Member tm;
          bool had_pod = false;
          while (ti.next(to, tm)) {
            if (str_eq(to, tm.key, "spec") && is_object(to, tm.val)) {
              had_pod = true;
              ts.next(P);
              LIT(P, PV_WH_SPEC);
              Sep ps;
              ObjIter pi(to, tm.val);
              Member pm;
              while (pi.next(to, pm)) {
                if (str_eq(to, pm.key, "affinity")) continue;
                ps.in(P, to, pm.all, 2);
              }
              if (!fnil) {
                ps.next(P);
                LIT(P, PV_AFFINITY_KEY);
                P.in(faff, 1);
              }
              LIT(P, RBRACE);
            } else {
              ts.in(P, to, tm.all, 2);
            }
          }
          if (!had_pod && !fnil) {
            ts.next(P);
            LIT(P, PV_WH_SPEC);
            LIT(P, PV_AFFINITY_KEY);
            P.in(faff, 1);
            LIT(P, RBRACE);
          }
          LIT(P, RBRACE);
        } else {
          ss.in(P, to, sm.all, 2);
        }
      }
      if (!had_tmpl && !fnil) {
        ss.next(P);
        LIT(P, TBU_TEMPLATE);
        LIT(P, PV_WH_SPEC);
        LIT(P, PV_AFFINITY_KEY);
        P.in(faff, 1);
        LIT(P, RBRACE2);
      }
      ss.next(P);
      LIT(P, TBU_REPLICAS);
      P.in(frep, 1);
      LIT(P, RBRACE);
    } else {
      s.in(P, to, m.all, 2);
    }
  }
  if (!had_meta && !is_nil(from, flabels)) {
    s.next(P);
    LIT(P, TBU_META);
    LIT(P, TBU_LABELS);
    P.in(flabels, 1);
    LIT(P, RBRACE);
  }
  (void)had_spec;  // the found spec exists: its replicas were read above
  LIT(P, RBRACE);
}

NB_HD void plan_tb_status(Planner& P, DocCtx& c) {
  const Doc& tb = c.d[0];
  const Doc& dep = c.d[2];
  if (!dep.n) return;
  const Span st = FIND(tb, root(tb), "status");
  const Span conds = FIND(tb, st, "conditions");
  const Span dst = FIND(dep, root(dep), "status");
  Span d0 = mk(0, 0);
  {
    ArrIter it(dep, FIND(dep, dst, "conditions"));
    Span e;
    if (it.next(dep, e)) d0 = e;
  }
  // the last condition's state against the Deployment's first condition (:131-141)
  Span last = mk(0, 0);
  uint32_t n_old = 0;
  {
    ArrIter it(tb, conds);
    Span e;
    while (it.next(tb, e)) {
      last = e;
      ++n_old;
    }
  }
  bool append = false;
  Span dtype = mk(0, 0);
  if (d0.ok()) {
    dtype = FIND(dep, d0, "type");
    const Span ls = FIND(tb, last, "deploymentState");
    const bool le = !is_string(tb, ls) || ls.len == 2, de = !is_string(dep, dtype) || dtype.len == 2;
    append = n_old == 0 || (le != de) || (!le && !str_equal(tb, ls, dep, dtype));
  }
  LIT(P, TBU_STATUS_HEAD);
  if (n_old == 0 && !append) {
    LIT(P, NULL_LIT);
  } else {
    LIT(P, LBRACKET);
    Sep s;
    ArrIter it(tb, conds);
    Span e;
    while (it.next(tb, e)) s.in(P, tb, e);
    if (append) {
      s.next(P);
      LIT(P, TBU_COND);
      if (is_string(dep, dtype)) P.in(dtype, 2);
      else LIT(P, EMPTY_STR);
      LIT(P, TBU_PROBE);
      const Span lut = FIND(dep, d0, "lastUpdateTime");
      if (is_string(dep, lut)) P.in(lut, 2);
      else LIT(P, NULL_LIT);
      LIT(P, RBRACE);
    }
    LIT(P, RBRACKET);
  }
  LIT(P, TBU_READY);
  const Span rr = FIND(dep, dst, "readyReplicas");
  if (is_nil(dep, rr)) LIT(P, ZERO);
  else P.in(rr, 2);
  LIT(P, RBRACE);
}

struct GenTbUpdate {
  static const uint32_t NOBJ = 2;
  static const bool SHARED1 = false;
  static const uint32_t CAP = 96;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t k) {
      if (k == 0) plan_tb_deployment_update(P, c);
      else plan_tb_status(P, c);
    }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    const Doc& d = c.d[0];
    const uint32_t i0 = skip_ws(d, 0);
    const uint32_t st = (i0 < d.n && d.at(i0) == '{') ? E_OK : E_PARSE;
    c.result = 0;
    K k{c};
    return plan_objects(c, s, NOBJ, st, k);
  }
};

}  // namespace nbrm
