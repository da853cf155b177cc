// nbrm_pd.cuh — PodDefault admission: select the PodDefaults that apply to a
// Pod, check that merging them is conflict-free, and emit the JSON patch.
//
// Reference: components/admission-webhook/main.go — mutatePods :599-704 (from
// the decoded Pod on), filterPodDefaults :72-97, safeToApplyPodDefaultsOnPod
// :101-166, the merge* family :170-475, applyPodDefaultsOnPod :480-556,
// applyPodDefaultsOnContainer :561-578, setCommandAndArgs :582-597.
// Label-selector semantics: k8s.io/apimachinery metav1.LabelSelectorAsSelector
// (not vendored).
//
// Input slot 0: the Pods of the admission requests, one per tuple.  Slot 1 is a
// SHARED arena of PodDefault objects; extra = two u32 per Pod, the [begin, end)
// range of that arena the reference would have listed for the Pod's namespace
// (null = the whole arena).  Output per Pod: an RFC 6902 patch
//   [{"op":"add","path":"/metadata","value":{..}},{"op":"add","path":"/spec","value":{..}}]
// (length 0 = allowed without a patch).  The reference's patch comes from
// mattbaird/jsonpatch with Go map iteration order, so it is compared by effect:
// applying it to the Pod gives the reference's mutated Pod.
// result = number of PodDefaults applied.
#pragma once
#include "nbrm_cmp.cuh"

namespace nbrm {


// Fixed room per Pod.  BASELINE.json's admission config lists up to 100 PodDefaults per
// namespace; the reference applies however many match (main.go:72-97 has no limit), so the
// room covers every one of them matching, with slack: a Pod beyond it is rejected with
// NBRM_E_LIMIT, never mutated partially.
static const uint32_t PD_MAX_MATCH = 128;  // PodDefaults applied to one Pod
static const uint32_t PD_MAX_ADDED = 256;  // items one merge may append to a list

NB_HD Doc pd_doc(const DocCtx& c, uint32_t j) {
  Doc d;
  const uint64_t o0 = c.shared_off[j], o1 = c.shared_off[j + 1];
  d.p = c.shared + o0;
  d.n = (uint32_t)(o1 - o0);
#ifdef NBRM_TAPE
  if (c.shared_tidx) {  // the PodDefault list stays resident and was indexed once: hop, do not scan
    const uint32_t* ti = reinterpret_cast<const uint32_t*>(c.shared_tidx) + 2 * (size_t)j;
    d.tape = reinterpret_cast<const TapeEnt*>(c.shared_tape) + ti[0];
    d.nt = ti[1] & 0xffffu;
    d.tflags = d.nt ? (ti[1] >> 16) : DOC_WS;
    d.dlim = TAPE_MAX_DEPTH;
  }
#endif
  return d;
}

// Out-of-line doors to the deep compare and the key lookup of nbrm_cmp.cuh: the
// merges below call them from many places, and one copy each keeps this kernel's
// compile time in seconds (inlined everywhere it took twenty minutes).
NB_FN bool pd_equal(const Doc& a, Span x, const Doc& b, Span y, bool* overflow) {
  return json_equal(a, x, b, y, overflow);
}
NB_FN Span pd_find_key(const Doc& d, Span obj, const Doc& kd, Span key) { return find_key(d, obj, kd, key); }

// String value or "" when the member is missing / not a string.
NB_FN bool strval_equal(const Doc& a, Span x, const Doc& b, Span y) {
  const bool xs = is_string(a, x) && x.len > 2, ys = is_string(b, y) && y.len > 2;
  if (!xs || !ys) return xs == ys;
  return str_equal(a, x, b, y);
}

template <uint32_t N>
NB_HD bool strval_is(const Doc& d, Span x, const char (&lit)[N]) {
  return str_eq(d, x, lit);
}

// 1 = the selector matches the labels, 0 = it does not, -1 = the selector is
// invalid (In/NotIn without values, Exists/DoesNotExist with values, unknown
// operator).  All requirements are validated before any is evaluated, as
// LabelSelectorAsSelector does.
NB_FN int selector_matches(const Doc& pd, Span sel, const Doc& pod, Span labels) {
  const Span exprs = find(pd, sel, "matchExpressions");
  {
    ArrIter it(pd, exprs);
    Span e;
    while (it.next(pd, e)) {
      const Span op = find(pd, e, "operator");
      const Span vals = find(pd, e, "values");
      const bool has_vals = is_array(pd, vals) && !is_empty_container(pd, vals);
      if (strval_is(pd, op, "In") || strval_is(pd, op, "NotIn")) {
        if (!has_vals) return -1;
      } else if (strval_is(pd, op, "Exists") || strval_is(pd, op, "DoesNotExist")) {
        if (has_vals) return -1;
      } else {
        return -1;
      }
    }
  }
  {
    ObjIter it(pd, find(pd, sel, "matchLabels"));
    Member m;
    while (it.next(pd, m)) {
      const Span pv = pd_find_key(pod, labels, pd, m.key);
      if (!is_string(pod, pv) || !is_string(pd, m.val) || !str_equal(pod, pv, pd, m.val)) return 0;
    }
  }
  ArrIter it(pd, exprs);
  Span e;
  while (it.next(pd, e)) {
    const Span key = find(pd, e, "key");
    const Span op = find(pd, e, "operator");
    Span pv = mk(0, 0);
    bool present = false;
    {
      ObjIter li(pod, labels);
      Member lm;
      while (li.next(pod, lm)) {
        // a label key is compared with the requirement's key ("" when missing)
        Doc kd = pod;
        Span kq = lm.key;
        if (strval_equal(kd, kq, pd, key)) {
          present = true;
          pv = lm.val;
        }
      }
    }
    bool in_vals = false;
    if (present) {
      ArrIter vi(pd, find(pd, e, "values"));
      Span v;
      while (vi.next(pd, v))
        if (is_string(pd, v) && is_string(pod, pv) && str_equal(pd, v, pod, pv)) in_vals = true;
    }
    if (strval_is(pd, op, "In")) {
      if (!present || !in_vals) return 0;
    } else if (strval_is(pd, op, "NotIn")) {
      if (present && in_vals) return 0;
    } else if (strval_is(pd, op, "Exists")) {
      if (!present) return 0;
    } else {
      if (present) return 0;
    }
  }
  return 1;
}

// One item appended by a merge: element (or map member) `s` of PodDefault `pd`.
struct Added {
  uint32_t pd;
  Span s;
};

struct AddList {
  Added a[PD_MAX_ADDED];
  uint32_t n;
};

// Everything a merge needs to know about the request.
struct PdCtx {
  const DocCtx* c;
  const Doc* pod;
  uint32_t match[PD_MAX_MATCH];  // indices into the shared arena, in list order
  uint32_t nmatch;
  bool conflict, overflow;
};

// merge* for a list keyed by one string field (main.go:170-213 imagePullSecrets,
// :217-252 env, :321-360 volumes, :364-407 containers, :411-450 tolerations):
// existing items keep their place, a PodDefault item whose key is new is
// appended, one whose key exists must be DeepEqual to the holder of that key.
// `existing` is the Pod's array (may be nil); field names the PodDefault spec
// list; key names the key field.
template <uint32_t NF, uint32_t NK>
NB_FN void merge_keyed(PdCtx& x, Span existing, const char (&field)[NF], const char (&key)[NK],
                       AddList& out) {
  out.n = 0;
  const Doc& pod = *x.pod;
  for (uint32_t q = 0; q < x.nmatch; ++q) {
    const Doc pd = pd_doc(*x.c, x.match[q]);
    ArrIter it(pd, find(pd, find(pd, root(pd), "spec"), field));
    Span v;
    while (it.next(pd, v)) {
      const Span vk = find(pd, v, key);
      // the holder of this key: the LAST existing item with it (Go map overwrite) ...
      Span held = mk(0, 0);
      {
        ArrIter ei(pod, existing);
        Span e;
        while (ei.next(pod, e))
          if (strval_equal(pod, find(pod, e, key), pd, vk)) held = e;
      }
      if (held.ok()) {
        if (!pd_equal(pod, held, pd, v, &x.overflow)) x.conflict = true;
        continue;
      }
      // ... or the first appended item with it
      bool found = false;
      for (uint32_t k = 0; k < out.n && !found; ++k) {
        const Doc od = pd_doc(*x.c, out.a[k].pd);
        if (strval_equal(od, find(od, out.a[k].s, key), pd, vk)) {
          found = true;
          if (!pd_equal(od, out.a[k].s, pd, v, &x.overflow)) x.conflict = true;
        }
      }
      if (found) continue;
      if (out.n >= PD_MAX_ADDED) {
        x.overflow = true;
        return;
      }
      out.a[out.n].pd = x.match[q];
      out.a[out.n].s = v;
      ++out.n;
    }
  }
}

// mergeVolumeMounts (:266-317): keyed by name for what is appended, and every
// PodDefault mount must also agree with whatever holds its mountPath.
NB_FN void merge_volume_mounts(PdCtx& x, Span existing, AddList& out, AddList& paths) {
  out.n = 0;
  paths.n = 0;
  const Doc& pod = *x.pod;
  for (uint32_t q = 0; q < x.nmatch; ++q) {
    const Doc pd = pd_doc(*x.c, x.match[q]);
    ArrIter it(pd, find(pd, find(pd, root(pd), "spec"), "volumeMounts"));
    Span v;
    while (it.next(pd, v)) {
      for (int pass = 0; pass < 2; ++pass) {  // 0: by name (:282-294), 1: by mountPath (:296-306)
        AddList& tab = pass ? paths : out;
        Span vk = pass ? find(pd, v, "mountPath") : find(pd, v, "name");
        Span held = mk(0, 0);
        ArrIter ei(pod, existing);
        Span e;
        while (ei.next(pod, e)) {
          const Span ek = pass ? find(pod, e, "mountPath") : find(pod, e, "name");
          if (strval_equal(pod, ek, pd, vk)) held = e;
        }
        if (held.ok()) {
          if (!pd_equal(pod, held, pd, v, &x.overflow)) x.conflict = true;
          continue;
        }
        bool found = false;
        for (uint32_t k = 0; k < tab.n && !found; ++k) {
          const Doc od = pd_doc(*x.c, tab.a[k].pd);
          const Span ok = pass ? find(od, tab.a[k].s, "mountPath") : find(od, tab.a[k].s, "name");
          if (strval_equal(od, ok, pd, vk)) {
            found = true;
            if (!pd_equal(od, tab.a[k].s, pd, v, &x.overflow)) x.conflict = true;
          }
        }
        if (found) continue;
        if (tab.n >= PD_MAX_ADDED) {
          x.overflow = true;
          return;
        }
        tab.a[tab.n].pd = x.match[q];
        tab.a[tab.n].s = v;
        ++tab.n;
      }
    }
  }
}

// mergeMap (:454-475) for annotations / labels: a PodDefault entry whose key is
// new is added, one whose key exists must carry the same value.  out holds the
// added members (key":"value spans of the PodDefault documents).
template <uint32_t NF>
NB_FN void merge_map(PdCtx& x, Span existing, const char (&field)[NF], AddList& out) {
  out.n = 0;
  const Doc& pod = *x.pod;
  for (uint32_t q = 0; q < x.nmatch; ++q) {
    const Doc pd = pd_doc(*x.c, x.match[q]);
    ObjIter it(pd, find(pd, find(pd, root(pd), "spec"), field));
    Member m;
    while (it.next(pd, m)) {
      const Span held = pd_find_key(pod, existing, pd, m.key);
      if (held.ok()) {
        if (!strval_equal(pod, held, pd, m.val)) x.conflict = true;
        continue;
      }
      bool found = false;
      for (uint32_t k = 0; k < out.n && !found; ++k) {
        const Doc od = pd_doc(*x.c, out.a[k].pd);
        // out.a[k].s spans key .. value: its key is the string at its start
        const uint32_t ke = skip_string(od, out.a[k].s.off);
        const Span okey = mk(out.a[k].s.off, ke - out.a[k].s.off);
        if (str_equal(od, okey, pd, m.key)) {
          found = true;
          uint32_t vs = skip_ws(od, ke);
          vs = skip_ws(od, vs + 1);  // past ':'
          const Span oval = mk(vs, out.a[k].s.end() - vs);
          if (!strval_equal(od, oval, pd, m.val)) x.conflict = true;
        }
      }
      if (found) continue;
      if (out.n >= PD_MAX_ADDED) {
        x.overflow = true;
        return;
      }
      out.a[out.n].pd = x.match[q];
      out.a[out.n].s = m.all;
      ++out.n;
    }
  }
}

// Segment from PodDefault j of the shared arena (input slot 1).
NB_HD void put_pd(Planner& P, const DocCtx& c, uint32_t j, Span s) {
  P.put(SRC_IN1, (uint32_t)c.shared_off[j] + s.off, s.len);
}

// "<key>":[ existing elements , appended items ]   — nothing when both are empty.
// Returns true when something was emitted.
NB_FN bool emit_merged_list(Planner& P, Sep& sep, const PdCtx& x, uint32_t key_off, uint32_t key_len,
                            Span existing, const AddList& add) {
  const Doc& pod = *x.pod;
  const bool have_existing = is_array(pod, existing) && !is_empty_container(pod, existing);
  if (!have_existing && !add.n) return false;
  sep.next(P);
  P.pool(key_off, key_len);
  LIT(P, LBRACKET);
  Sep es;
  if (have_existing) {
    P.in(mk(existing.off + 1, existing.len - 2));
    es.first = false;
  }
  for (uint32_t k = 0; k < add.n; ++k) {
    es.next(P);
    put_pd(P, *x.c, add.a[k].pd, add.a[k].s);
  }
  LIT(P, RBRACKET);
  return true;
}
#define EMIT_LIST(P, sep, x, name, existing, add) \
  emit_merged_list(P, sep, x, ::nbrm::L_##name, ::nbrm::L_##name##_END - ::nbrm::L_##name, existing, add)

// metadata.name of PodDefault j (string value with quotes, or nothing).
NB_HD Span pd_name(const Doc& pd) {
  return find(pd, find(pd, root(pd), "metadata"), "name");
}

// The Pod's annotation key `key` is the marker of one of the applied PodDefaults
// ("poddefault.admission.kubeflow.org/poddefault-<name>", :553-555).
NB_FN bool is_marker_key(const PdCtx& x, Span key) {
  const Doc& pod = *x.pod;
  const Span k = inner(key);
  const char prefix[] = "poddefault.admission.kubeflow.org/poddefault-";
  const uint32_t np = sizeof(prefix) - 1;
  if (k.len < np) return false;
  for (uint32_t i = 0; i < np; ++i)
    if (pod.at(k.off + i) != prefix[i]) return false;
  const Span rest = mk(k.off + np, k.len - np);
  for (uint32_t q = 0; q < x.nmatch; ++q) {
    const Doc pd = pd_doc(*x.c, x.match[q]);
    const Span nm = pd_name(pd);
    const Span ni = is_string(pd, nm) ? inner(nm) : mk(0, 0);
    if (spans_eq(pod, rest, pd, ni)) return true;
  }
  return false;
}

// One container with env / volumeMounts / envFrom merged and command / args
// defaulted (applyPodDefaultsOnContainer :561-578, setCommandAndArgs :582-597).
NB_FN void plan_pd_container(Planner& P, PdCtx& x, Span ctr, AddList& la, AddList& lb) {
  const Doc& pod = *x.pod;
  const DocCtx& c = *x.c;
  const bool istio = str_eq(pod, find(pod, ctr, "name"), "istio-proxy");
  // command / args: the first applied PodDefault that has one, when the container has none
  int cmd_pd = -1, args_pd = -1;
  Span cmd_s = mk(0, 0), args_s = mk(0, 0);
  if (!istio) {
    const bool has_cmd = !is_nil(pod, find(pod, ctr, "command"));
    const bool has_args = !is_nil(pod, find(pod, ctr, "args"));
    for (uint32_t q = 0; q < x.nmatch; ++q) {
      const Doc pd = pd_doc(c, x.match[q]);
      const Span sp = find(pd, root(pd), "spec");
      if (!has_cmd && cmd_pd < 0) {
        const Span v = find(pd, sp, "command");
        if (!is_nil(pd, v)) {
          cmd_pd = (int)x.match[q];
          cmd_s = v;
        }
      }
      if (!has_args && args_pd < 0) {
        const Span v = find(pd, sp, "args");
        if (!is_nil(pd, v)) {
          args_pd = (int)x.match[q];
          args_s = v;
        }
      }
    }
  }
  LIT(P, LBRACE);
  Sep s;
  bool had_env = false, had_vm = false, had_ef = false;
  ObjIter it(pod, ctr);
  Member m;
  while (it.next(pod, m)) {
    if (KEY_IS(pod, m, "env")) {
      had_env = true;
      merge_keyed(x, m.val, "env", "name", la);
      EMIT_LIST(P, s, x, KEY_ENV, m.val, la);
    } else if (KEY_IS(pod, m, "volumeMounts")) {
      had_vm = true;
      merge_volume_mounts(x, m.val, la, lb);
      EMIT_LIST(P, s, x, KEY_VOLUMEMOUNTS, m.val, la);
    } else if (KEY_IS(pod, m, "envFrom")) {
      had_ef = true;
      la.n = 0;  // mergeEnvFrom (:254-262): plain concatenation
      for (uint32_t q = 0; q < x.nmatch && !x.overflow; ++q) {
        const Doc pd = pd_doc(c, x.match[q]);
        ArrIter ei(pd, find(pd, find(pd, root(pd), "spec"), "envFrom"));
        Span e;
        while (ei.next(pd, e)) {
          if (la.n >= PD_MAX_ADDED) {
            x.overflow = true;
            break;
          }
          la.a[la.n].pd = x.match[q];
          la.a[la.n].s = e;
          ++la.n;
        }
      }
      EMIT_LIST(P, s, x, KEY_ENVFROM, m.val, la);
    } else if ((KEY_IS(pod, m, "command") && cmd_pd >= 0) || (KEY_IS(pod, m, "args") && args_pd >= 0)) {
      continue;  // a null member: replaced below
    } else {
      s.in(P, pod, m.all);
    }
  }
  if (!had_env) {
    merge_keyed(x, mk(0, 0), "env", "name", la);
    EMIT_LIST(P, s, x, KEY_ENV, mk(0, 0), la);
  }
  if (!had_vm) {
    merge_volume_mounts(x, mk(0, 0), la, lb);
    EMIT_LIST(P, s, x, KEY_VOLUMEMOUNTS, mk(0, 0), la);
  }
  if (!had_ef) {
    la.n = 0;
    for (uint32_t q = 0; q < x.nmatch && !x.overflow; ++q) {
      const Doc pd = pd_doc(c, x.match[q]);
      ArrIter ei(pd, find(pd, find(pd, root(pd), "spec"), "envFrom"));
      Span e;
      while (ei.next(pd, e)) {
        if (la.n >= PD_MAX_ADDED) {
          x.overflow = true;
          break;
        }
        la.a[la.n].pd = x.match[q];
        la.a[la.n].s = e;
        ++la.n;
      }
    }
    EMIT_LIST(P, s, x, KEY_ENVFROM, mk(0, 0), la);
  }
  if (cmd_pd >= 0) {
    s.next(P);
    LIT(P, KEY_COMMAND);
    put_pd(P, c, (uint32_t)cmd_pd, cmd_s);
  }
  if (args_pd >= 0) {
    s.next(P);
    LIT(P, KEY_ARGS);
    put_pd(P, c, (uint32_t)args_pd, args_s);
  }
  LIT(P, RBRACE);
}

NB_HD void plan_poddefault(Planner& P, DocCtx& c) {
  const Doc& pod = c.d[0];
  const Span rootv = root(pod);
  const Span meta = FIND(pod, rootv, "metadata");
  const Span spec = FIND(pod, rootv, "spec");
  const Span labels = FIND(pod, meta, "labels");
  const Span annotations = FIND(pod, meta, "annotations");
  const Span ns = FIND(pod, meta, "namespace");
  if (is_object(pod, annotations)) {  // :625-633
    if (str_eq(pod, find(pod, annotations, "poddefault.admission.kubeflow.org/exclude"), "true")) return;
    if (find(pod, annotations, "kubernetes.io/config.mirror").ok()) return;
  }
  uint32_t begin = 0, end = c.shared_n;
  if (c.extra) {
    begin = ((const uint32_t*)c.extra)[2 * c.index];
    end = ((const uint32_t*)c.extra)[2 * c.index + 1];
    if (end > c.shared_n) end = c.shared_n;
  }
  PdCtx x;
  x.c = &c;
  x.pod = &pod;
  x.nmatch = 0;
  x.conflict = x.overflow = false;
  for (uint32_t j = begin; j < end; ++j) {  // filterPodDefaults :72-97
    const Doc pd = pd_doc(c, j);
    const Span pr = root(pd);
    const int r = selector_matches(pd, find(pd, find(pd, pr, "spec"), "selector"), pod, labels);
    if (r < 0) {
      P.err = E_SELECTOR;
      return;
    }
    if (!r) continue;
    if (!strval_equal(pd, find(pd, find(pd, pr, "metadata"), "namespace"), pod, ns)) continue;  // :87
    if (x.nmatch >= PD_MAX_MATCH) {
      P.err = E_LIMIT;
      return;
    }
    x.match[x.nmatch++] = j;
  }
  if (!x.nmatch) return;
  c.result = x.nmatch;

  AddList la, lb;
  LIT(P, PATCH_HEAD);
  {  // metadata: annotations and labels merged (:519-529), markers added (:553-555)
    LIT(P, LBRACE);
    Sep ms;
    ObjIter it(pod, meta);
    Member m;
    while (it.next(pod, m)) {
      if (KEY_IS(pod, m, "annotations") || KEY_IS(pod, m, "labels")) continue;
      ms.in(P, pod, m.all);
    }
    merge_map(x, annotations, "annotations", la);
    ms.next(P);
    LIT(P, KEY_ANNOTATIONS);
    LIT(P, LBRACE);
    Sep as;
    {
      ObjIter ai(pod, annotations);
      Member am;
      while (ai.next(pod, am)) {
        if (is_marker_key(x, am.key)) continue;  // reassigned below
        as.in(P, pod, am.all);
      }
    }
    for (uint32_t k = 0; k < la.n; ++k) {
      as.next(P);
      put_pd(P, c, la.a[k].pd, la.a[k].s);
    }
    for (uint32_t q = 0; q < x.nmatch; ++q) {
      const Doc pd = pd_doc(c, x.match[q]);
      const Span nm = pd_name(pd);
      const Span ni = is_string(pd, nm) ? inner(nm) : mk(0, 0);
      bool later = false;  // two PodDefaults of one name write one map key: the last wins
      for (uint32_t q2 = q + 1; q2 < x.nmatch && !later; ++q2) {
        const Doc p2 = pd_doc(c, x.match[q2]);
        const Span n2 = pd_name(p2);
        later = spans_eq(pd, ni, p2, is_string(p2, n2) ? inner(n2) : mk(0, 0));
      }
      if (later) continue;
      as.next(P);
      LIT(P, MARKER_PREFIX);
      put_pd(P, c, x.match[q], ni);
      LIT(P, QUOTE_COLON);
      const Span rv = find(pd, find(pd, root(pd), "metadata"), "resourceVersion");
      if (is_string(pd, rv)) put_pd(P, c, x.match[q], rv);
      else LIT(P, EMPTY_STR);
    }
    LIT(P, RBRACE);
    merge_map(x, labels, "labels", la);
    if ((is_object(pod, labels) && !is_empty_container(pod, labels)) || la.n) {
      ms.next(P);
      LIT(P, KEY_LABELS);
      LIT(P, LBRACE);
      Sep ls;
      ObjIter li(pod, labels);
      Member lm;
      while (li.next(pod, lm)) ls.in(P, pod, lm.all);
      for (uint32_t k = 0; k < la.n; ++k) {
        ls.next(P);
        put_pd(P, c, la.a[k].pd, la.a[k].s);
      }
      LIT(P, RBRACE);
    }
    LIT(P, RBRACE);
  }
  LIT(P, RBRACE);  // closes the first operation
  LIT(P, PATCH_MID);
  {
    // serviceAccountName / automountServiceAccountToken: the last PodDefault that sets one (:512-517)
    int sa_pd = -1, am_pd = -1;
    Span sa_s = mk(0, 0), am_s = mk(0, 0);
    for (uint32_t q = 0; q < x.nmatch; ++q) {
      const Doc pd = pd_doc(c, x.match[q]);
      const Span sp = find(pd, root(pd), "spec");
      const Span a = find(pd, sp, "automountServiceAccountToken");
      if (!is_nil(pd, a)) {
        am_pd = (int)x.match[q];
        am_s = a;
      }
      const Span sa = find(pd, sp, "serviceAccountName");
      if (is_string(pd, sa) && sa.len > 2) {
        sa_pd = (int)x.match[q];
        sa_s = sa;
      }
    }
    const Span containers = FIND(pod, spec, "containers");
    LIT(P, LBRACE);
    Sep ss;
    bool had_vol = false, had_tol = false, had_ips = false, had_init = false, had_ctr = false;
    ObjIter it(pod, spec);
    Member m;
    while (it.next(pod, m)) {
      if (KEY_IS(pod, m, "volumes")) {
        had_vol = true;
        merge_keyed(x, m.val, "volumes", "name", la);
        EMIT_LIST(P, ss, x, KEY_VOLUMES, m.val, la);
      } else if (KEY_IS(pod, m, "tolerations")) {
        had_tol = true;
        merge_keyed(x, m.val, "tolerations", "key", la);
        EMIT_LIST(P, ss, x, KEY_TOLERATIONS, m.val, la);
      } else if (KEY_IS(pod, m, "imagePullSecrets")) {
        had_ips = true;
        merge_keyed(x, m.val, "imagePullSecrets", "name", la);
        EMIT_LIST(P, ss, x, KEY_IMAGEPULLSECRETS, m.val, la);
      } else if (KEY_IS(pod, m, "initContainers")) {
        had_init = true;
        merge_keyed(x, m.val, "initContainers", "name", la);
        EMIT_LIST(P, ss, x, KEY_INITCONTAINERS, m.val, la);
      } else if (KEY_IS(pod, m, "containers")) {
        had_ctr = true;
        // every container gets env / mounts / envFrom first; sidecars are appended
        // after that and so receive none of it (:531-550)
        const bool any = is_array(pod, m.val) && !is_empty_container(pod, m.val);
        AddList side;
        merge_keyed(x, m.val, "sidecars", "name", side);
        if (any || side.n) {
          ss.next(P);
          LIT(P, KEY_CONTAINERS_ARR);
          LIT(P, LBRACKET);
          Sep cs;
          ArrIter ci(pod, m.val);
          Span e;
          while (ci.next(pod, e)) {
            cs.next(P);
            if (is_object(pod, e)) plan_pd_container(P, x, e, la, lb);
            else P.in(e);
          }
          for (uint32_t k = 0; k < side.n; ++k) {
            cs.next(P);
            put_pd(P, c, side.a[k].pd, side.a[k].s);
          }
          LIT(P, RBRACKET);
        }
      } else if (KEY_IS(pod, m, "serviceAccountName") && sa_pd >= 0) {
        continue;
      } else if (KEY_IS(pod, m, "automountServiceAccountToken") && am_pd >= 0) {
        continue;
      } else {
        ss.in(P, pod, m.all);
      }
    }
    if (!had_vol) {
      merge_keyed(x, mk(0, 0), "volumes", "name", la);
      EMIT_LIST(P, ss, x, KEY_VOLUMES, mk(0, 0), la);
    }
    if (!had_tol) {
      merge_keyed(x, mk(0, 0), "tolerations", "key", la);
      EMIT_LIST(P, ss, x, KEY_TOLERATIONS, mk(0, 0), la);
    }
    if (!had_ips) {
      merge_keyed(x, mk(0, 0), "imagePullSecrets", "name", la);
      EMIT_LIST(P, ss, x, KEY_IMAGEPULLSECRETS, mk(0, 0), la);
    }
    if (!had_init) {
      merge_keyed(x, mk(0, 0), "initContainers", "name", la);
      EMIT_LIST(P, ss, x, KEY_INITCONTAINERS, mk(0, 0), la);
    }
    if (!had_ctr) {
      merge_keyed(x, mk(0, 0), "sidecars", "name", la);
      EMIT_LIST(P, ss, x, KEY_CONTAINERS_ARR, mk(0, 0), la);
    }
    (void)containers;
    if (sa_pd >= 0) {
      ss.next(P);
      LIT(P, KEY_SERVICEACCOUNTNAME);
      put_pd(P, c, (uint32_t)sa_pd, sa_s);
    }
    if (am_pd >= 0) {
      ss.next(P);
      LIT(P, KEY_AUTOMOUNT);
      put_pd(P, c, (uint32_t)am_pd, am_s);
    }
    LIT(P, RBRACE);
  }
  LIT(P, PATCH_TAIL);
  if (x.overflow) P.err = E_LIMIT;
  else if (x.conflict) P.err = E_CONFLICT;  // :671-679
}

struct GenPodDefault {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = true;  // slot 1 = the PodDefault arena, shared by every Pod
  static const uint32_t CAP = 4096;  // 128 PodDefaults x (volume + mount per container + marker + ...) segments
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { plan_poddefault(P, c); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    K k{c};
    uint32_t st = is_object(c.d[0], root(c.d[0])) ? E_OK : E_PARSE;
    return plan_objects(c, s, NOBJ, st, k);
  }
};

}  // namespace nbrm
