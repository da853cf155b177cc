// nbrm_patch.cuh — RFC 6902 JSON patches and the AdmissionReview envelope.
//
// Reference:
//   odh-notebook-controller/controllers/notebook_webhook.go:296-301 — Handle answers with
//     admission.PatchResponseFromRaw(req.Object.Raw, marshaledNotebook): the JSON patch that
//     turns the submitted Notebook into the mutated one (gomodules.xyz/jsonpatch/v2, not
//     vendored), wrapped by controller-runtime in an AdmissionReview;
//   admission-webhook/main.go:685-701 — jsonpatch.CreatePatch(podCopyJSON, podJSON)
//     (mattbaird/jsonpatch), and :708-749 serve: the AdmissionReview around it, uid echoed,
//     patch base64-encoded by encoding/json ([]byte), patchType "JSONPatch".
// Neither library's operation order is deterministic (Go map iteration), so patches are
// compared by effect: applying the patch to the original gives the mutated document.
//
// GenJsonDiff  slot 0 = original, slot 1 = mutated (an empty mutated document: the stage
//              before rejected it, nothing is emitted).  Output: the patch, length 0 when the
//              two are equal (the reference then sends no patch and no patchType).
//              Objects are diffed member by member (add / remove / replace, recursing into
//              objects that both sides have), arrays that only grew get one "add .../-" per
//              new element, arrays of equal length are diffed element by element, any other
//              change replaces the value.  Values are copied from the mutated document.
// GenReviewOpen   slot 0 = AdmissionReview requests.  Output: request.object (with
//              metadata.namespace filled from request.namespace when the object has none,
//              admission-webhook/main.go:617-620), or request.oldObject with flag REVIEW_OLD.
//              result word: bit 0 the resource is core/v1 pods, bit 1 operation CREATE,
//              bit 2 operation UPDATE.
// GenReviewClose  slot 0 = AdmissionReview requests, slot 1 = the patches (or empty), extra =
//              one u32 per review, the status word of the stage that produced the patch.
//              Output: the AdmissionReview response.  flags bit 0: controller-runtime style
//              (kind / apiVersion echoed, denial with code 403) for the Notebook webhook;
//              clear: admission-webhook style ({"response":{...}}, errors as status.message).
#pragma once
#include "nbrm_cmp.cuh"
#include "nbrm_odh.cuh"  // b64c

namespace nbrm {

static const int DIFF_MAX_DEPTH = 12;

// One level of the path being diffed: a member key (span of the mutated or original
// document) or an array index.
struct PathElem {
  Span key;       // with quotes; len 0: an index
  uint32_t doc;   // which document the key's bytes are in (input slot)
  uint32_t index;
};

struct DiffCtx {
  Planner* P;
  DocCtx* c;
  const Doc* a;  // original (slot 0)
  const Doc* b;  // mutated  (slot 1)
  PathElem path[DIFF_MAX_DEPTH + 1];
  int depth;
  Sep ops;
  bool deep;     // a compare ran out of depth: no answer
};

// A key as a JSON-pointer token: '~' -> "~0", '/' -> "~1" (RFC 6901); the JSON escapes of
// the key stay as they are, the pointer lives inside a JSON string again.
NB_FN void put_pointer_token(DiffCtx& x, const Doc& d, Span key, uint32_t slot) {
  const Span k = inner(key);
  bool plain = true;
  for (uint32_t i = 0; i < k.len; ++i) {
    const char ch = d.at(k.off + i);
    if (ch == '~' || ch == '/') plain = false;
  }
  Planner& P = *x.P;
  if (plain) {
    P.in(k, slot);
    return;
  }
  Aux& aux = x.c->aux;
  if (aux.used + 2 * k.len > aux.cap) {
    P.err = E_LIMIT;
    return;
  }
  char* o = aux.p + aux.used;
  uint32_t n = 0;
  for (uint32_t i = 0; i < k.len; ++i) {
    const char ch = d.at(k.off + i);
    if (ch == '~') {
      o[n++] = '~';
      o[n++] = '0';
    } else if (ch == '/') {
      o[n++] = '~';
      o[n++] = '1';
    } else {
      o[n++] = ch;
    }
  }
  P.aux(aux.abs + aux.used, n);
  aux.used += n;
}

NB_FN void put_index(DiffCtx& x, uint32_t v) {
  Aux& aux = x.c->aux;
  Planner& P = *x.P;
  if (aux.used + 12 > aux.cap) {
    P.err = E_LIMIT;
    return;
  }
  char tmp[12];
  uint32_t n = 0;
  do {
    tmp[n++] = (char)('0' + v % 10);
    v /= 10;
  } while (v);
  char* o = aux.p + aux.used;
  for (uint32_t i = 0; i < n; ++i) o[i] = tmp[n - 1 - i];
  P.aux(aux.abs + aux.used, n);
  aux.used += n;
}

// {"op":"<op>","path":"<path so far>[/<last>]"[,"value":<value from the mutated document>]}
// last_kind: 0 nothing more, 1 a key (last_key in document last_slot), 2 "-" (append), 3 an index
NB_FN void emit_op(DiffCtx& x, int op, int last_kind, Span last_key, uint32_t last_slot, uint32_t last_index,
                   Span value) {
  Planner& P = *x.P;
  x.ops.next(P);
  if (op == 0) LIT(P, PATCH_OP_ADD);
  else if (op == 1) LIT(P, PATCH_OP_REMOVE);
  else LIT(P, PATCH_OP_REPLACE);
  for (int k = 0; k < x.depth; ++k) {
    LIT(P, SLASH);
    if (x.path[k].key.len) put_pointer_token(x, x.path[k].doc ? *x.b : *x.a, x.path[k].key, x.path[k].doc);
    else put_index(x, x.path[k].index);
  }
  if (last_kind) LIT(P, SLASH);
  if (last_kind == 1) put_pointer_token(x, last_slot ? *x.b : *x.a, last_key, last_slot);
  else if (last_kind == 2) LIT(P, DASH);
  else if (last_kind == 3) put_index(x, last_index);
  if (op == 1) {
    LIT(P, PATCH_OP_END_NOVALUE);
  } else {
    LIT(P, PATCH_OP_VALUE);
    P.in(value, 1);
    LIT(P, RBRACE);
  }
}

struct DiffFrame {
  ObjIter ib;        // objects: members of the mutated side
  ObjIter ia;        // objects, second phase: members of the original side (removals)
  ArrIter ea, eb;    // arrays of equal length
  Span va, vb;       // the two containers
  uint32_t index;    // arrays: position of the next pair
  uint8_t is_obj, phase;
};

// Diff of two values that both exist; the path to them is x.path[0..depth).
NB_FN void diff_values(DiffCtx& x, Span va0, Span vb0) {
  const Doc& a = *x.a;
  const Doc& b = *x.b;
  DiffFrame st[DIFF_MAX_DEPTH];
  int sp = 0;
  const int base_depth = x.depth;
  // push the first pair (the caller made sure both are containers of one kind, or calls replace itself)
  Span va = va0, vb = vb0;
  bool push = true;
  for (;;) {
    if (push) {
      x.depth = base_depth + sp;  // the path now ends at this container
      DiffFrame& f = st[sp];
      f.va = va;
      f.vb = vb;
      f.index = 0;
      f.phase = 0;
      f.is_obj = is_object(b, vb) ? 1 : 0;
      if (f.is_obj) {
        f.ib = ObjIter(b, vb);
        f.ia = ObjIter(a, va);
      } else {
        // arrays: how long, and how far do they agree
        uint32_t na = 0, nb = 0, same = 0;
        {
          ArrIter ia(a, va), ib(b, vb);
          Span ea, eb;
          bool ga = ia.next(a, ea), gb = ib.next(b, eb), agree = true;
          while (ga || gb) {
            if (ga && gb && agree && json_equal(a, ea, b, eb, &x.deep)) ++same;
            else agree = false;
            na += ga;
            nb += gb;
            if (ga) ga = ia.next(a, ea);
            if (gb) gb = ib.next(b, eb);
          }
        }
        if (same == na && nb > na) {  // only grew: append the new elements
          ArrIter ib(b, vb);
          Span eb;
          uint32_t k = 0;
          while (ib.next(b, eb)) {
            if (k++ < na) continue;
            emit_op(x, 0, 2, mk(0, 0), 0, 0, eb);
          }
          f.phase = 9;  // done
        } else if (na != nb) {  // anything else that changes the length: the whole array
          emit_op(x, 2, 0, mk(0, 0), 0, 0, vb);
          f.phase = 9;
        } else {
          f.ea = ArrIter(a, va);
          f.eb = ArrIter(b, vb);
        }
      }
      ++sp;
      push = false;
    }
    if (sp == 0) break;
    DiffFrame& f = st[sp - 1];
    x.depth = base_depth + sp - 1;
    if (f.phase == 9) {
      --sp;
      continue;
    }
    if (f.is_obj) {
      if (f.phase == 0) {
        Member m;
        if (!f.ib.next(b, m)) {
          f.phase = 1;
          continue;
        }
        if (is_nil(b, m.val)) continue;  // null = absent (Go nil)
        const Span av = find_key(a, f.va, b, m.key);
        if (is_nil(a, av)) {
          emit_op(x, 0, 1, m.key, 1, 0, m.val);
        } else if (!json_equal(a, av, b, m.val, &x.deep)) {
          const bool both_obj = is_object(a, av) && is_object(b, m.val);
          const bool both_arr = is_array(a, av) && is_array(b, m.val);
          if ((both_obj || both_arr) && sp < DIFF_MAX_DEPTH && x.depth < DIFF_MAX_DEPTH) {
            x.path[x.depth].key = m.key;
            x.path[x.depth].doc = 1;
            va = av;
            vb = m.val;
            push = true;
          } else {
            emit_op(x, 2, 1, m.key, 1, 0, m.val);
          }
        }
      } else {
        Member m;
        if (!f.ia.next(a, m)) {
          --sp;
          continue;
        }
        if (is_nil(a, m.val)) continue;
        if (is_nil(b, find_key(b, f.vb, a, m.key))) emit_op(x, 1, 1, m.key, 0, 0, mk(0, 0));
      }
    } else {
      Span ea, eb;
      if (!f.ea.next(a, ea) || !f.eb.next(b, eb)) {
        --sp;
        continue;
      }
      const uint32_t idx = f.index++;
      if (json_equal(a, ea, b, eb, &x.deep)) continue;
      const bool both_obj = is_object(a, ea) && is_object(b, eb);
      if (both_obj && sp < DIFF_MAX_DEPTH && x.depth < DIFF_MAX_DEPTH) {
        x.path[x.depth].key = mk(0, 0);
        x.path[x.depth].index = idx;
        va = ea;
        vb = eb;
        push = true;
      } else {
        emit_op(x, 2, 3, mk(0, 0), 0, idx, eb);
      }
    }
  }
  x.depth = base_depth;
}

NB_HD void plan_json_diff(Planner& P, DocCtx& c) {
  const Doc& a = c.d[0];
  const Doc& b = c.d[1];
  const Span ra = root(a), rb = root(b);
  if (!is_object(a, ra) || !is_object(b, rb)) {
    P.err = E_PARSE;
    return;
  }
  bool deep = false;
  if (json_equal(a, ra, b, rb, &deep)) return;  // nothing to patch
  if (deep) {
    P.err = E_LIMIT;
    return;
  }
  DiffCtx x;
  x.P = &P;
  x.c = &c;
  x.a = &a;
  x.b = &b;
  x.depth = 0;
  x.deep = false;
  LIT(P, LBRACKET);
  diff_values(x, ra, rb);
  LIT(P, RBRACKET);
  if (x.deep && !P.err) P.err = E_LIMIT;
}

struct GenJsonDiff {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;
  static const uint32_t CAP = 512;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { plan_json_diff(P, c); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    if (c.d[1].n == 0) {  // the stage before produced nothing for this request
      s.nseg[0] = 0;
      s.lens[0] = 0;
      return E_OK;
    }
    K k{c};
    return plan_objects(c, s, NOBJ, E_OK, k);
  }
};

// ---------------------------------------------------------------- AdmissionReview
enum ReviewBits : uint32_t { REVIEW_IS_POD = 1u, REVIEW_CREATE = 2u, REVIEW_UPDATE = 4u };
enum ReviewFlags : uint32_t { REVIEW_CTRL_RUNTIME = 1u, REVIEW_OLD = 2u, REVIEW_PODS_ONLY = 4u };

NB_HD bool review_is_pod(const Doc& d, Span req) {
  const Span res = FIND(d, req, "resource");
  const Span grp = FIND(d, res, "group");
  return (!is_string(d, grp) || grp.len == 2) && str_eq(d, FIND(d, res, "version"), "v1") &&
         str_eq(d, FIND(d, res, "resource"), "pods");
}

NB_HD void plan_review_open(Planner& P, DocCtx& c) {
  const Doc& d = c.d[0];
  const Span req = FIND(d, root(d), "request");
  if (!is_object(d, req)) {
    P.err = E_PARSE;
    return;
  }
  if (c.flags & REVIEW_OLD) {
    const Span old = FIND(d, req, "oldObject");
    if (is_object(d, old)) P.in(old);
    return;
  }
  uint32_t bits = 0;
  if (review_is_pod(d, req)) bits |= REVIEW_IS_POD;
  const Span op = FIND(d, req, "operation");
  if (str_eq(d, op, "CREATE")) bits |= REVIEW_CREATE;
  if (str_eq(d, op, "UPDATE")) bits |= REVIEW_UPDATE;
  c.result = bits;
  const Span obj = FIND(d, req, "object");
  if (!is_object(d, obj)) {
    P.err = E_PARSE;
    return;
  }
  // pod.Namespace = ar.Request.Namespace when the manifest has none (admission-webhook/main.go:617-620)
  const Span rns = FIND(d, req, "namespace");
  const Span meta = FIND(d, obj, "metadata");
  const Span ons = FIND(d, meta, "namespace");
  const bool fill = is_string(d, rns) && rns.len > 2 && (!is_string(d, ons) || ons.len == 2);
  if (!fill) {
    P.in(obj);
    return;
  }
  LIT(P, LBRACE);
  Sep s;
  bool had_meta = false;
  ObjIter it(d, obj);
  Member m;
  while (it.next(d, m)) {
    if (KEY_IS(d, m, "metadata") && is_object(d, m.val)) {
      had_meta = true;
      s.next(P);
      LIT(P, KEY_METADATA);
      LIT(P, LBRACE);
      Sep ms;
      ObjIter mi(d, m.val);
      Member mm;
      while (mi.next(d, mm)) {
        if (KEY_IS(d, mm, "namespace")) continue;
        ms.in(P, d, mm.all);
      }
      ms.next(P);
      LIT(P, KEY_NAMESPACE);
      P.in(rns);
      LIT(P, RBRACE);
    } else if (KEY_IS(d, m, "metadata")) {
      continue;  // null metadata: replaced below
    } else {
      s.in(P, d, m.all);
    }
  }
  if (!had_meta) {
    s.next(P);
    LIT(P, KEY_METADATA);
    LIT(P, LBRACE);
    LIT(P, KEY_NAMESPACE);
    P.in(rns);
    LIT(P, RBRACE);
  }
  LIT(P, RBRACE);
}

struct GenReviewOpen {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;
  static const uint32_t CAP = 64;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { plan_review_open(P, c); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    K k{c};
    const uint32_t st = is_object(c.d[0], root(c.d[0])) ? E_OK : E_PARSE;
    return plan_objects(c, s, NOBJ, st, k);
  }
};

// base64 of the bytes of `s` (document d) into the aux arena; returns false when it does not fit.
NB_FN bool aux_base64(DocCtx& c, const Doc& d, Span s, uint32_t* off, uint32_t* len) {
  Aux& aux = c.aux;
  const uint32_t need = (s.len + 2) / 3 * 4;
  if (aux.used + need > aux.cap) return false;
  char* o = aux.p + aux.used;
  uint32_t n = 0, i = 0;
  for (; i + 3 <= s.len; i += 3) {
    const uint32_t w = ((uint32_t)(unsigned char)d.at(s.off + i) << 16) |
                       ((uint32_t)(unsigned char)d.at(s.off + i + 1) << 8) | (unsigned char)d.at(s.off + i + 2);
    o[n++] = b64c(w >> 18);
    o[n++] = b64c((w >> 12) & 63);
    o[n++] = b64c((w >> 6) & 63);
    o[n++] = b64c(w & 63);
  }
  if (s.len - i == 1) {
    const uint32_t w = (uint32_t)(unsigned char)d.at(s.off + i) << 16;
    o[n++] = b64c(w >> 18);
    o[n++] = b64c((w >> 12) & 63);
    o[n++] = '=';
    o[n++] = '=';
  } else if (s.len - i == 2) {
    const uint32_t w = ((uint32_t)(unsigned char)d.at(s.off + i) << 16) |
                       ((uint32_t)(unsigned char)d.at(s.off + i + 1) << 8);
    o[n++] = b64c(w >> 18);
    o[n++] = b64c((w >> 12) & 63);
    o[n++] = b64c((w >> 6) & 63);
    o[n++] = '=';
  }
  *off = aux.abs + aux.used;
  *len = n;
  aux.used += n;
  return true;
}

NB_HD void plan_review_close(Planner& P, DocCtx& c) {
  const Doc& d = c.d[0];
  const Doc& patch = c.d[1];
  const Span rt = root(d);
  const Span req = FIND(d, rt, "request");
  const Span uid = FIND(d, req, "uid");
  const uint32_t st = c.extra ? ((const uint32_t*)c.extra)[c.index] : 0u;
  const bool ctrl = (c.flags & REVIEW_CTRL_RUNTIME) != 0;
  LIT(P, LBRACE);
  if ((c.flags & REVIEW_PODS_ONLY) && !review_is_pod(d, req)) {
    // mutatePods returns nil for anything but core/v1 pods and serve then writes a review
    // without a response (admission-webhook/main.go:602-605, :733-737)
    LIT(P, RBRACE);
    return;
  }
  if (ctrl) {  // controller-runtime answers in the review version it was asked in
    const Span kind = FIND(d, rt, "kind"), api = FIND(d, rt, "apiVersion");
    LIT(P, KEY_KIND);
    if (is_string(d, kind)) P.in(kind); else LIT(P, REVIEW_KIND_DEFAULT);
    LIT(P, COMMA);
    LIT(P, KEY_APIVERSION);
    if (is_string(d, api)) P.in(api); else LIT(P, REVIEW_API_DEFAULT);
    LIT(P, COMMA);
  }
  LIT(P, REVIEW_RESPONSE_UID);  // "response":{"uid":
  if (is_string(d, uid)) P.in(uid); else LIT(P, EMPTY_STR);
  if (st == E_OK) {
    LIT(P, REVIEW_ALLOWED_TRUE);
    if (patch.n) {
      uint32_t off = 0, len = 0;
      if (!aux_base64(c, patch, mk(0, patch.n), &off, &len)) {
        P.err = E_LIMIT;
        return;
      }
      LIT(P, REVIEW_PATCH_OPEN);  // the fields in the order of the Go struct: patch, then patchType
      P.aux(off, len);
      LIT(P, REVIEW_PATCH_CLOSE);
    }
  } else if (ctrl) {
    if (st == E_DENIED) LIT(P, REVIEW_DENIED_BOTH);   // admission.Denied (notebook_webhook.go:273-274)
    else LIT(P, REVIEW_ERRORED_400);                  // admission.Errored(http.StatusBadRequest, err) (:244)
  } else {
    if (st == E_CONFLICT) LIT(P, REVIEW_STATUS_CONFLICT);       // main.go:671-679 via toAdmissionResponse
    else if (st == E_SELECTOR) LIT(P, REVIEW_STATUS_SELECTOR);  // :659-662
    else LIT(P, REVIEW_STATUS_ERROR);
  }
  LIT(P, RBRACE);
  LIT(P, RBRACE);
}

struct GenReviewClose {
  static const uint32_t NOBJ = 1;
  static const bool SHARED1 = false;
  static const uint32_t CAP = 32;
  struct K {
    DocCtx& c;
    NB_HD void operator()(Planner& P, uint32_t) { plan_review_close(P, c); }
  };
  NB_HD static uint32_t run(DocCtx& c, ObjSink s) {
    K k{c};
    const uint32_t st = is_object(c.d[0], root(c.d[0])) ? E_OK : E_PARSE;
    return plan_objects(c, s, NOBJ, st, k);
  }
};

}  // namespace nbrm
