// nbrm_plan.cuh — the plan an object generator writes and the emit kernel reads.
//
// Every generator is split in two:
//   plan  (one input tuple per thread, branchy, reads the input JSON once)
//         -> a short list of segments per output object + its byte length
//   emit  (one output object per thread, lanes of a warp on the same object
//          kind; pure gather-copy)
//         -> packed output bytes at the offsets an exclusive scan of the
//            lengths gives
// A segment is (source buffer, offset, length).  Sources are the input arenas
// (up to three per call: e.g. Notebook, found StatefulSet, Pod), the literal
// pool (template text, built once per context) and the per-call aux arena
// (bytes a planner had to compute: unescaped text, formatted numbers/times).
#pragma once
#include <stdint.h>

#include "nbrm_json.cuh"

namespace nbrm {

enum Src : uint32_t { SRC_IN0 = 0, SRC_IN1 = 1, SRC_IN2 = 2, SRC_POOL = 3, SRC_AUX = 4, SRC_COUNT = 5 };

enum Err : uint32_t {
  E_OK = 0,
  E_PARSE = 1,          // not a JSON object / a required object is missing
  E_NO_CONTAINERS = 2,  // reference would panic: Containers[0] (notebook_controller.go:455,489)
  E_SEGS = 3,           // plan longer than the segment capacity
  E_PORTS_EMPTY = 4,    // reference would panic: Ports[0] on "ports":[] (notebook_controller.go:491)
  E_NIL_REPLICAS = 5,   // reference would panic: *Spec.Replicas (reconcilehelper/util.go:123)
  E_DENIED = 6,         // admission denied by the reference's own rule
  E_CONFLICT = 7,       // PodDefault merge conflict (admission-webhook/main.go:671-679)
  E_LIMIT = 8,          // a documented fixed limit of this implementation was exceeded
  E_SELECTOR = 9,       // a PodDefault's label selector does not convert (admission-webhook/main.go:77-79)
  E_ENV = 10,           // an environment variable the reference requires is not set, or has a value it rejects
};

struct Seg {
  uint32_t off;
  uint32_t len_src;  // len in the low 28 bits, Src in the high 4
};

static const uint32_t SEG_LEN_MASK = 0x0fffffffu;

struct Planner {
  Seg* segs;
  uint32_t cap;
  uint32_t n;      // segments already written to segs[]
  uint32_t total;
  uint32_t err;
  uint32_t base[3];  // absolute offset of this thread's document inside input arena k
  // the segment being built lives in registers until the next one starts:
  // consecutive pieces of the same source merge without touching memory
  uint32_t cur_off, cur_len, cur_src;

  NB_HD void begin(Seg* s, uint32_t c) {
    segs = s;
    cap = c;
    n = 0;
    total = 0;
    cur_len = 0;
    cur_off = 0;
    cur_src = 0;
  }
  NB_HD void flush() {
    if (!cur_len) return;
    if (n < cap) {
      segs[n].off = cur_off;
      segs[n].len_src = (cur_src << 28) | cur_len;
      ++n;
    } else {
      err = E_SEGS;
    }
    cur_len = 0;
  }
  NB_HD void put(uint32_t src, uint32_t off, uint32_t len) {
    if (!len) return;
    total += len;
    if (cur_len && cur_src == src && cur_off + cur_len == off && cur_len + len <= SEG_LEN_MASK) {
      cur_len += len;
      return;
    }
    flush();
    while (len > SEG_LEN_MASK) {  // a span of 256 MiB or more goes out as several segments
      cur_src = src;
      cur_off = off;
      cur_len = SEG_LEN_MASK;
      flush();
      off += SEG_LEN_MASK;
      len -= SEG_LEN_MASK;
    }
    cur_src = src;
    cur_off = off;
    cur_len = len;
  }
  // number of segments once the open one is written
  NB_HD uint32_t finish() {
    flush();
    return n;
  }
  NB_HD void in(Span s, uint32_t k = 0) { put(k, base[k] + s.off, s.len); }
  NB_HD void pool(uint32_t off, uint32_t len) { put(SRC_POOL, off, len); }
  NB_HD void aux(uint32_t off, uint32_t len) { put(SRC_AUX, off, len); }
};

// ---------------------------------------------------------------- literal pool
// NBRM_LITS(X) lists every template fragment.  The enum chain gives each one a
// fixed offset in the pool (one spare byte between neighbours); the host fills
// the pool once per context from the same list.
#include "nbrm_lits.inc"

enum LitOff : uint32_t {
  L__BEGIN = 0,
#define X(name, text) L_##name, L_##name##_END = L_##name + sizeof(text) - 1,
  NBRM_LITS(X)
#undef X
      L__TOTAL
};

#define LIT(P, name) (P).pool(::nbrm::L_##name, ::nbrm::L_##name##_END - ::nbrm::L_##name)

// The fragments' text at compile time (LT_name), for planners that must know where
// something lies inside a fragment they emit (views, below).
#define X(name, text) static constexpr char LT_##name[] = text;
NBRM_LITS(X)
#undef X

// Index just after the first occurrence of `needle` in `text` (compile time).
template <uint32_t N, uint32_t M>
constexpr uint32_t lit_after(const char (&text)[N], const char (&needle)[M]) {
  for (uint32_t i = 0; i + (M - 1) <= N - 1; ++i) {
    bool hit = true;
    for (uint32_t k = 0; k + 1 < M; ++k)
      if (text[i + k] != needle[k]) {
        hit = false;
        break;
      }
    if (hit) return i + M - 1;
  }
  return 0xffffffffu;
}
#define LIT_AFTER(name, needle) (::nbrm::lit_after(::nbrm::LT_##name, needle))
#define LIT_LEN(name) ((uint32_t)sizeof(::nbrm::LT_##name) - 1u)

// ---------------------------------------------------------------- views
// A generator knows where it puts things.  For the objects a later stage compares
// (notebook-controller's three, the ODH objects), it writes down where the parts that
// stage looks for lie in the object it planned — exactly the spans view_object
// (nbrm_cmp.cuh) would find by walking the emitted text — so the compare stage of the same
// pass starts from them instead of parsing what the library itself just wrote.  A slot
// filled by nbrm_upload has no views and is walked as before.
enum ViewField : uint32_t {
  VF_ROOT, VF_KIND, VF_META, VF_LABELS, VF_ANNOTATIONS, VF_SPEC, VF_REPLICAS, VF_TLABELS, VF_TPOD,
  VF_SELECTOR, VF_PORTS, VF_SUBJECTS, VIEW_FIELDS
};
struct ViewRec {
  uint32_t off[VIEW_FIELDS];
  uint32_t len[VIEW_FIELDS];
};
// The views of a batch are kept field-major: (off, len) of field f of object j at
// p[f * n + j].  Threads that plan or read neighbouring objects then touch neighbouring
// words, where an array of records would put every lane in a cache line of its own.  The
// table is zeroed before the plan kernel runs; a planner stores only the fields it has.
struct alignas(8) ViewEnt {
  uint32_t off, len;
};
struct ViewTable {
  ViewEnt* p;  // VIEW_FIELDS * n entries, or null
  uint32_t n;  // objects in the batch
  NB_HD void store(uint32_t j, uint32_t f, uint32_t off, uint32_t len) const {
    ViewEnt e;
    e.off = off;
    e.len = len;
    p[(size_t)f * n + j] = e;
  }
  NB_HD Span get(uint32_t j, uint32_t f) const {
    const ViewEnt e = p[(size_t)f * n + j];
    return mk(e.off, e.len);
  }
};
// What a planner writes through: one object's column of the table.
struct ViewOut {
  ViewTable t;  // t.p null: nobody asked
  uint32_t j;
  uint32_t b[VIEW_FIELDS];  // where a field began (begin ... end); constant indices, so these are registers
  NB_HD void clear() {}
  NB_HD void set(uint32_t f, uint32_t off, uint32_t end) {
    if (t.p) t.store(j, f, off, end - off);
  }
  NB_HD void begin(uint32_t f, uint32_t off) { b[f] = off; }
  NB_HD void end(uint32_t f, uint32_t e) {
    if (t.p) t.store(j, f, b[f], e - b[f]);
  }
};
// The views of one input document, as a planner receives them.
struct ViewIn {
  ViewTable t;
  uint32_t j;
  NB_HD bool ok() const { return t.p != nullptr; }
  NB_HD Span get(uint32_t f) const { return t.get(j, f); }
};

// Strings that come from the reference's process environment / flags.  The
// host JSON-escapes them and appends them to the pool after L__TOTAL.
// A piece of the literal pool.  Not a Span: this struct travels inside the kernel
// arguments between units compiled with and without the tape (a Span then carries a tape
// index), so its layout must not depend on that.
struct PoolSpan {
  uint32_t off, len;
};
struct EnvLits {
  PoolSpan istio_host;      // ISTIO_HOST or "*"            (notebook_controller.go:548-551)
  PoolSpan istio_gateway;   // ISTIO_GATEWAY or default     (:557-560)
  PoolSpan cluster_domain;  // CLUSTER_DOMAIN or default    (:537-539)
  PoolSpan ctrl_namespace;  // ODH controller namespace     (odh main.go:63-72)
  PoolSpan oauth_image;     // --oauth-proxy-image          (odh main.go:80)
  // profile-controller (profile_controller.go:419-504, :127-137)
  PoolSpan prof_when_key;   // "request.headers[<userid header>]" (quoted)
  PoolSpan prof_prefix;     // userid prefix (unquoted, escaped)
  PoolSpan prof_igw, prof_kfp, prof_nbc;  // principals (quoted)
  PoolSpan prof_ns_labels;  // members of the Namespace's labels map, rendered by the host
  // tensorboard-controller (tensorboard_controller.go:172,:325,:330): as set, no defaults (quoted)
  PoolSpan tb_image, tb_gateway, tb_host;
  // pvcviewer webhook: the defaults file's PodSpec (pvcviewer_webhook.go:78-91), split by the host
  PoolSpan pv_default_rest, pv_default_volumes;
  uint32_t add_fsgroup; // ADD_FSGROUP unset or "true"  (notebook_controller.go:475)
  uint32_t use_istio;   // USE_ISTIO == "true"          (:213)
  uint32_t pv_default_set;  // a defaults file with a non-zero PodSpec was given
  uint32_t tb_env;      // TbEnvBits (nbrm_tb.cuh): which of those are set; RWO_PVC_SCHEDULING on / invalid (:471-482)
};

// Separator-aware writer for the members of one JSON object or array.
struct Sep {
  bool first;
  NB_HD Sep() : first(true) {}
  NB_HD void next(Planner& P) {
    if (!first) LIT(P, COMMA);
    first = false;
  }
  // A member/element copied from input arena k.  When it directly follows the
  // previous copied bytes and the input already has the comma, the comma is
  // taken from the input so the run stays one segment.
  NB_HD void in(Planner& P, const Doc& d, Span s, uint32_t k = 0) {
    if (!first && P.cur_len && P.cur_src == k && s.off > 0 &&
        P.cur_off + P.cur_len == P.base[k] + s.off - 1 && d.at(s.off - 1) == ',') {
      P.cur_len += 1 + s.len;
      P.total += 1 + s.len;
      return;
    }
    next(P);
    P.in(s, k);
  }
};

}  // namespace nbrm
