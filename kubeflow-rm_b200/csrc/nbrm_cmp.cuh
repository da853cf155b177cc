// nbrm_cmp.cuh — "does the found object need an update, and what is written
// back": the reconcilehelper Copy*Fields functions on JSON text.
//
// Reference: components/common/reconcilehelper/util.go — CopyStatefulSetFields
// :107-134, CopyServiceFields :166-195, CopyVirtualService :199-219; the
// template-label pre-copy of notebook_controller.go:165-169; and the
// create-or-update decision around them (notebook_controller.go:147-179,
// :189-210, :636-657).
//
// Input slot 0 holds the desired objects ("from"), slot 1 the found objects
// ("to", an empty document = NotFound), tuple i pairs document i of both.
// Output: one object per pair —
//   found missing            -> the desired object as is   (caller Creates it)
//   update required          -> the found object with the owned fields copied
//                               over                       (caller Updates it)
//   nothing to do            -> length 0
#pragma once
#include "nbrm_gen.cuh"

namespace nbrm {

// ---------------------------------------------------------------- DeepEqual on JSON text
// reflect.DeepEqual of two decoded values, restated on the encoded text:
// objects compare by key (order-free), a member whose value is null counts as
// absent (Go nil), arrays compare element by element, strings compare decoded
// (escapes and raw UTF-8 alike), numbers and literals compare as tokens.

// Decoding cursor over the inside of a JSON string: yields Unicode code points
// (escapes and raw UTF-8 both decoded), -1 at the end.
struct StrCur {
  uint32_t i, end;
  NB_HD static int hex4(const Doc& d, uint32_t at, uint32_t end) {
    if (at + 4 > end) return -1;
    int v = 0;
    for (int k = 0; k < 4; ++k) {
      char h = d.at(at + k);
      int hv = (h >= '0' && h <= '9') ? h - '0'
               : (h >= 'a' && h <= 'f') ? h - 'a' + 10
               : (h >= 'A' && h <= 'F') ? h - 'A' + 10 : -1;
      if (hv < 0) return -1;
      v = v * 16 + hv;
    }
    return v;
  }
  NB_HD int next(const Doc& d) {
    if (i >= end) return -1;
    unsigned char c = (unsigned char)d.at(i);
    if (c == '\\') {
      if (i + 1 >= end) {
        i = end;
        return 0xFFFD;
      }
      char x = d.at(i + 1);
      i += 2;
      switch (x) {
        case 'b': return '\b';
        case 'f': return '\f';
        case 'n': return '\n';
        case 'r': return '\r';
        case 't': return '\t';
        case 'u': {
          int v = hex4(d, i, end);
          if (v < 0) {
            i = end;
            return 0xFFFD;
          }
          i += 4;
          if (v >= 0xD800 && v < 0xDC00 && i + 6 <= end && d.at(i) == '\\' && d.at(i + 1) == 'u') {
            int lo = hex4(d, i + 2, end);
            if (lo >= 0xDC00 && lo < 0xE000) {
              i += 6;
              return 0x10000 + ((v - 0xD800) << 10) + (lo - 0xDC00);
            }
          }
          if (v >= 0xD800 && v < 0xE000) return 0xFFFD;
          return v;
        }
        default: return (unsigned char)x;  // \" \\ \/
      }
    }
    if (c < 0x80) {
      ++i;
      return c;
    }
    int need = (c >= 0xF0) ? 3 : (c >= 0xE0) ? 2 : (c >= 0xC0) ? 1 : 0;
    int v = (need == 3) ? (c & 0x07) : (need == 2) ? (c & 0x0F) : (need == 1) ? (c & 0x1F) : 0xFFFD;
    ++i;
    for (int k = 0; k < need && i < end; ++k, ++i) v = (v << 6) | ((unsigned char)d.at(i) & 0x3F);
    return v;
  }
};

NB_HD bool str_equal(const Doc& a, Span x, const Doc& b, Span y) {
  if (spans_eq(a, x, b, y)) return true;
  StrCur ca, cb;
  Span ix = inner(x), iy = inner(y);
  ca.i = ix.off;
  ca.end = ix.end();
  cb.i = iy.off;
  cb.end = iy.end();
  for (;;) {
    int u = ca.next(a), v = cb.next(b);
    if (u != v) return false;
    if (u == -1) return true;
  }
}

// A member that decodes to Go nil.  SEM adds equality.Semantic.DeepEqual's rule
// that an empty slice or map equals a nil one (odh-notebook-controller/
// controllers/notebook_webhook.go:353,359).
template <bool SEM>
NB_HD bool member_is_nil(const Doc& d, Span v) {
  if (is_nil(d, v)) return true;
  if (SEM && (is_object(d, v) || is_array(d, v)) && is_empty_container(d, v)) return true;
  return false;
}

template <bool SEM>
NB_HD uint32_t count_members(const Doc& d, Span obj) {  // nil-valued members do not count
  ObjIter it(d, obj);
  Member m;
  uint32_t n = 0;
  while (it.next(d, m))
    if (!member_is_nil<SEM>(d, m.val)) ++n;
  return n;
}

// obj[key] where key is a span of another document (decoded compare).
NB_HD Span find_key(const Doc& d, Span obj, const Doc& kd, Span key) {
  ObjIter it(d, obj);
  Member m;
  while (it.next(d, m))
    if (str_equal(d, m.key, kd, key)) return m.val;
  return mk(0, 0);
}

static const int EQ_MAX_DEPTH = 24;

// One open container of the compare: where A's walk stands, and B's side of it.
struct EqFrame {
  ObjIter oa, ob;  // objects: A's members; B's members for the "same order" fast path
  ArrIter aa, ab;  // arrays: both element walks
  Span bobj;       // objects: B's whole object, for the key search
  uint32_t is_obj;
};

// nil (absent / null) on both sides is equal; nil vs anything else is not.
// *overflow is set when the values nest deeper than EQ_MAX_DEPTH: the answer (false) is
// then not a finding, and the callers report E_LIMIT instead of planning an update.
template <bool SEM>
NB_HD bool json_equal_t(const Doc& a, Span x, const Doc& b, Span y, bool* overflow = nullptr) {
  const bool xn = member_is_nil<SEM>(a, x), yn = member_is_nil<SEM>(b, y);
  if (xn || yn) return xn && yn;
  if (spans_eq(a, x, b, y)) return true;  // same text: the common steady-state case, one linear pass
  EqFrame st[EQ_MAX_DEPTH];
  int sp = 0;
  // compare one pair of values; containers are pushed
  Span cx = x, cy = y;
  bool have = true;
  for (;;) {
    if (have) {
      char ta = a.at(cx.off), tb = b.at(cy.off);
      if (ta == '"') {
        if (tb != '"' || !str_equal(a, cx, b, cy)) return false;
      } else if (ta == '{') {
        if (tb != '{' || count_members<SEM>(a, cx) != count_members<SEM>(b, cy)) return false;
        if (sp >= EQ_MAX_DEPTH) {
          if (overflow) *overflow = true;
          return false;
        }
        st[sp].oa = ObjIter(a, cx);
        st[sp].ob = ObjIter(b, cy);
        st[sp].bobj = cy;
        st[sp].is_obj = 1;
        ++sp;
      } else if (ta == '[') {
        if (tb != '[') return false;
        if (sp >= EQ_MAX_DEPTH) {
          if (overflow) *overflow = true;
          return false;
        }
        st[sp].aa = ArrIter(a, cx);
        st[sp].ab = ArrIter(b, cy);
        st[sp].is_obj = 0;
        ++sp;
      } else {
        if (!spans_eq(a, cx, b, cy)) return false;
      }
      have = false;
    }
    if (sp == 0) return true;
    EqFrame& f = st[sp - 1];
    if (f.is_obj) {
      Member m;
      bool got = false;
      while (f.oa.next(a, m)) {
        if (member_is_nil<SEM>(a, m.val)) continue;
        got = true;
        break;
      }
      if (!got) {
        --sp;
        continue;
      }
      // objects written by the same encoder list their members in the same order:
      // try B's next member before searching B for the key
      Span bv = mk(0, 0);
      {
        ObjIter ib = f.ob;
        Member bm;
        if (ib.next(b, bm) && str_equal(b, bm.key, a, m.key)) {
          bv = bm.val;
          f.ob = ib;
        }
      }
      if (!bv.ok()) bv = find_key(b, f.bobj, a, m.key);
      if (member_is_nil<SEM>(b, bv)) return false;
      cx = m.val;
      cy = bv;
      have = true;
    } else {
      Span ea, eb;
      bool ga = f.aa.next(a, ea), gb = f.ab.next(b, eb);
      if (ga != gb) return false;
      if (!ga) {
        --sp;
        continue;
      }
      cx = ea;
      cy = eb;
      have = true;
    }
  }
}

NB_HD bool json_equal(const Doc& a, Span x, const Doc& b, Span y, bool* overflow = nullptr) {
  return json_equal_t<false>(a, x, b, y, overflow);
}
NB_HD bool semantic_equal(const Doc& a, Span x, const Doc& b, Span y, bool* overflow = nullptr) {
  return json_equal_t<true>(a, x, b, y, overflow);
}

// ---------------------------------------------------------------- Copy*Fields

// util.go:109-121: for k, v := range to.<field> { if from.<field>[k] != v { requireUpdate } }
NB_HD bool map_field_differs(const Doc& from, Span fmap, const Doc& to, Span tmap) {
  ObjIter it(to, tmap);
  Member m;
  while (it.next(to, m)) {
    if (!is_string(to, m.val)) continue;
    Span fv = find_key(from, fmap, to, m.key);
    if (is_string(from, fv)) {
      if (!str_equal(from, fv, to, m.val)) return true;
    } else if (m.val.len > 2) {
      return true;  // missing in from reads as "" there
    }
  }
  return false;
}

// Spans of an owned object, found in one descent through the document.
struct ObjView {
  Span root, kind, meta, labels, annotations, spec;
  Span replicas, tmpl, tmeta, tlabels, tpod;  // StatefulSet: spec.replicas, spec.template.{metadata{labels},spec}
  Span selector, ports;                       // Service: spec.selector, spec.ports
};

NB_HD ObjView view_object(const Doc& d) {
  ObjView v;
  v.root = v.kind = v.meta = v.labels = v.annotations = v.spec = mk(0, 0);
  v.replicas = v.tmpl = v.tmeta = v.tlabels = v.tpod = v.selector = v.ports = mk(0, 0);
  Cursor c;
  c.i = skip_ws(d, 0);
  const uint32_t rs = c.i;
  if (!obj_enter(d, c)) return v;
  Member m;
  while (obj_key(d, c, m)) {
    if (KEY_IS(d, m, "kind")) {
      v.kind = val_skip(d, c);
    } else if (KEY_IS(d, m, "metadata")) {
      const uint32_t s0 = c.i;
      if (!obj_enter(d, c)) {
        val_skip(d, c);
        continue;
      }
      Member mm;
      while (obj_key(d, c, mm)) {
        const Span val = val_skip(d, c);
        if (KEY_IS(d, mm, "labels")) v.labels = val;
        else if (KEY_IS(d, mm, "annotations")) v.annotations = val;
      }
      v.meta = mk(s0, c.i - s0);
    } else if (KEY_IS(d, m, "spec")) {
      const uint32_t s0 = c.i;
      if (!obj_enter(d, c)) {
        v.spec = val_skip(d, c);
        continue;
      }
      Member sm;
      while (obj_key(d, c, sm)) {
        if (KEY_IS(d, sm, "template")) {
          const uint32_t t0 = c.i;
          if (!obj_enter(d, c)) {
            v.tmpl = val_skip(d, c);
            continue;
          }
          Member tm;
          while (obj_key(d, c, tm)) {
            if (KEY_IS(d, tm, "metadata")) {
              const uint32_t m0 = c.i;
              if (!obj_enter(d, c)) {
                v.tmeta = val_skip(d, c);
                continue;
              }
              Member lm;
              while (obj_key(d, c, lm)) {
                const Span val = val_skip(d, c);
                if (KEY_IS(d, lm, "labels")) v.tlabels = val;
              }
              v.tmeta = mk(m0, c.i - m0);
            } else {
              const Span val = val_skip(d, c);
              if (KEY_IS(d, tm, "spec")) v.tpod = val;
            }
          }
          v.tmpl = mk(t0, c.i - t0);
        } else {
          const Span val = val_skip(d, c);
          if (KEY_IS(d, sm, "replicas")) v.replicas = val;
          else if (KEY_IS(d, sm, "selector")) v.selector = val;
          else if (KEY_IS(d, sm, "ports")) v.ports = val;
        }
      }
      v.spec = mk(s0, c.i - s0);
    } else {
      val_skip(d, c);
    }
  }
  v.root = mk(rs, c.i - rs);
  return v;
}

NB_HD void plan_merged_metadata(Planner& P, const Doc& from, const ObjView& fv, const Doc& to,
                                const ObjView& tv, const Member& meta_member) {
  P.in(mk(meta_member.key.off, meta_member.val.off - meta_member.key.off), 1);  // "metadata":
  LIT(P, LBRACE);
  Sep s;
  ObjIter it(to, tv.meta);
  Member m;
  while (it.next(to, m)) {
    if (KEY_IS(to, m, "labels") || KEY_IS(to, m, "annotations")) continue;
    s.in(P, to, m.all, 1);
  }
  if (!is_nil(from, fv.labels)) {  // to.Labels = from.Labels (:114)
    s.next(P);
    LIT(P, KEY_LABELS);
    P.in(fv.labels, 0);
  }
  if (!is_nil(from, fv.annotations)) {  // :121
    s.next(P);
    LIT(P, KEY_ANNOTATIONS);
    P.in(fv.annotations, 0);
  }
  LIT(P, RBRACE);
}

// The spans a generator wrote down for an object it planned (nbrm_plan.cuh, views): what
// view_object finds, without the walk.  Fields the generators do not record stay empty; the
// compare below asks a desired object for none of them.
NB_HD ObjView view_from_table(const ViewIn& r) {
  ObjView v;
  v.root = r.get(VF_ROOT);
  v.kind = r.get(VF_KIND);
  v.meta = r.get(VF_META);
  v.labels = r.get(VF_LABELS);
  v.annotations = r.get(VF_ANNOTATIONS);
  v.spec = r.get(VF_SPEC);
  v.replicas = r.get(VF_REPLICAS);
  v.tlabels = r.get(VF_TLABELS);
  v.tpod = r.get(VF_TPOD);
  v.selector = r.get(VF_SELECTOR);
  v.ports = r.get(VF_PORTS);
  v.tmpl = v.tmeta = mk(0, 0);
  return v;
}

template <uint32_t N>
NB_HD bool kind_is(const Doc& d, Span rootv, const char (&k)[N]) {
  return str_eq(d, FIND(d, rootv, "kind"), k);
}

// The planner for one (desired, found) pair.
NB_HD void plan_copy_fields(Planner& P, const Doc& from, const Doc& to, const ViewIn& from_view) {
  if (to.n == 0 || to.at(skip_ws(to, 0)) != '{') {  // NotFound -> Create(desired)
    P.in(from_view.ok() ? from_view.get(VF_ROOT) : root(from), 0);
    return;
  }
  ObjView fv = from_view.ok() ? view_from_table(from_view) : view_object(from), tv = view_object(to);
  if (!is_object(from, fv.root)) {
    P.err = E_PARSE;
    return;
  }
  bool deep = false;  // set when a compare met nesting beyond EQ_MAX_DEPTH
  bool require = map_field_differs(from, fv.labels, to, tv.labels);        // util.go:109-113
  require |= map_field_differs(from, fv.annotations, to, tv.annotations);  // :116-120

  if (str_eq(from, fv.kind, "VirtualService")) {  // CopyVirtualService, util.go:199-219
    if (!is_object(from, fv.spec)) return;         // :200-206 from has no spec map: false
    bool upd = !is_object(to, tv.spec) || !json_equal(from, fv.spec, to, tv.spec, &deep);
    if (deep) P.err = E_LIMIT;  // nested deeper than the compare follows: no answer, not "differs"
    if (!upd || deep) return;
    LIT(P, LBRACE);
    Sep s;
    ObjIter it(to, tv.root);
    Member m;
    while (it.next(to, m)) {
      if (KEY_IS(to, m, "spec")) continue;
      s.in(P, to, m.all, 1);
    }
    s.next(P);
    LIT(P, KEY_SPEC);
    P.in(fv.spec, 0);
    LIT(P, RBRACE);
    return;
  }

  if (str_eq(from, fv.kind, "Service")) {  // CopyServiceFields, util.go:166-195
    Span fsel = fv.selector, fports = fv.ports;
    Span tsel = tv.selector, tports = tv.ports;
    if (!json_equal(from, fsel, to, tsel, &deep)) require = true;      // :184
    if (!json_equal(from, fports, to, tports, &deep)) require = true;  // :189
    if (deep) P.err = E_LIMIT;
    if (!require || deep) return;
    LIT(P, LBRACE);
    Sep s;
    bool had_spec = false;
    ObjIter it(to, tv.root);
    Member m;
    while (it.next(to, m)) {
      if (KEY_IS(to, m, "metadata")) {
        s.next(P);
        plan_merged_metadata(P, from, fv, to, tv, m);
      } else if (KEY_IS(to, m, "spec") && is_object(to, m.val)) {
        had_spec = true;
        s.next(P);
        LIT(P, KEY_SPEC);
        LIT(P, LBRACE);
        Sep ss;
        ObjIter si(to, m.val);
        Member sm;
        while (si.next(to, sm)) {
          if (KEY_IS(to, sm, "selector") || KEY_IS(to, sm, "ports")) continue;
          ss.in(P, to, sm.all, 1);
        }
        if (!is_nil(from, fsel)) {
          ss.next(P);
          LIT(P, KEY_SELECTOR);
          P.in(fsel, 0);
        }
        if (!is_nil(from, fports)) {
          ss.next(P);
          LIT(P, KEY_PORTS);
          P.in(fports, 0);
        }
        LIT(P, RBRACE);
      } else {
        s.in(P, to, m.all, 1);
      }
    }
    if (!had_spec) {
      s.next(P);
      LIT(P, KEY_SPEC);
      LIT(P, LBRACE);
      Sep ss;
      if (!is_nil(from, fsel)) {
        ss.next(P);
        LIT(P, KEY_SELECTOR);
        P.in(fsel, 0);
      }
      if (!is_nil(from, fports)) {
        ss.next(P);
        LIT(P, KEY_PORTS);
        P.in(fports, 0);
      }
      LIT(P, RBRACE);
    }
    LIT(P, RBRACE);
    return;
  }

  // StatefulSet: notebook_controller.go:165-169 then CopyStatefulSetFields, util.go:107-134
  Span frep = fv.replicas, trep = tv.replicas;
  if (is_nil(from, frep) || is_nil(to, trep)) {
    P.err = E_NIL_REPLICAS;  // util.go:123 dereferences both
    return;
  }
  Span ttmpl = tv.tmpl;
  Span ftlabels = fv.tlabels, ttlabels = tv.tlabels;
  Span fpod = fv.tpod, tpod = tv.tpod;
  const bool replicas_differ = !spans_eq(from, frep, to, trep);
  // pre-copy: template labels follow only when the replica counts differ
  const bool copy_tlabels = replicas_differ && !json_equal(from, ftlabels, to, ttlabels, &deep);
  if (replicas_differ) require = true;                             // :123-126
  if (!json_equal(from, fpod, to, tpod, &deep)) require = true;    // :128-130
  if (deep) P.err = E_LIMIT;
  if (!require || deep) return;

  LIT(P, LBRACE);
  Sep s;
  ObjIter it(to, tv.root);
  Member m;
  while (it.next(to, m)) {
    if (KEY_IS(to, m, "metadata")) {
      s.next(P);
      plan_merged_metadata(P, from, fv, to, tv, m);
    } else if (KEY_IS(to, m, "spec")) {
      s.next(P);
      LIT(P, KEY_SPEC);
      LIT(P, LBRACE);
      Sep ss;
      ObjIter si(to, m.val);
      Member sm;
      while (si.next(to, sm)) {
        if (KEY_IS(to, sm, "replicas")) {
          ss.next(P);
          LIT(P, KEY_REPLICAS);
          P.in(frep, 0);
        } else if (KEY_IS(to, sm, "template")) {
          ss.next(P);
          LIT(P, KEY_TEMPLATE);
          LIT(P, LBRACE);
          Sep ts;
          bool had_meta = false;
          ObjIter ti(to, sm.val);
          Member tm;
          while (ti.next(to, tm)) {
            if (KEY_IS(to, tm, "spec")) continue;  // replaced below
            if (KEY_IS(to, tm, "metadata") && copy_tlabels) {
              had_meta = true;
              ts.next(P);
              LIT(P, KEY_METADATA);
              LIT(P, LBRACE);
              Sep ms;
              ObjIter mi(to, tm.val);
              Member mm;
              while (mi.next(to, mm)) {
                if (KEY_IS(to, mm, "labels")) continue;
                ms.in(P, to, mm.all, 1);
              }
              if (!is_nil(from, ftlabels)) {
                ms.next(P);
                LIT(P, KEY_LABELS);
                P.in(ftlabels, 0);
              }
              LIT(P, RBRACE);
              continue;
            }
            if (KEY_IS(to, tm, "metadata")) had_meta = true;
            ts.in(P, to, tm.all, 1);
          }
          if (!had_meta && copy_tlabels && !is_nil(from, ftlabels)) {
            ts.next(P);
            LIT(P, KEY_METADATA);
            LIT(P, LBRACE);
            LIT(P, KEY_LABELS);
            P.in(ftlabels, 0);
            LIT(P, RBRACE);
          }
          if (!is_nil(from, fpod)) {  // to.Spec.Template.Spec = from.Spec.Template.Spec (:131)
            ts.next(P);
            LIT(P, KEY_SPEC);
            P.in(fpod, 0);
          }
          LIT(P, RBRACE);
        } else {
          ss.in(P, to, sm.all, 1);
        }
      }
      if (!ttmpl.ok() && !is_nil(from, fpod)) {  // found had no template at all
        ss.next(P);
        LIT(P, KEY_TEMPLATE);
        LIT(P, LBRACE);
        LIT(P, KEY_SPEC);
        P.in(fpod, 0);
        LIT(P, RBRACE);
      }
      LIT(P, RBRACE);
    } else {
      s.in(P, to, m.all, 1);
    }
  }
  LIT(P, RBRACE);
}

// The compare generator: slot 0 desired, slot 1 found, one output per pair.
struct GenCopyFields {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;  // slot 1 is per-tuple, not one shared arena
  static const uint32_t CAP = 48;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { plan_copy_fields(P, c.d[0], c.d[1], c.view[0]); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    K k{c};
    uint32_t st = E_OK;
    if (c.d[0].n == 0) {  // nothing desired (the generator rejected the Notebook): nothing to do
      s.nseg[0] = 0;
      s.lens[0] = 0;
      return E_OK;
    }
    return plan_objects(c, s, NOBJ, st, k);
  }
};

}  // namespace nbrm
