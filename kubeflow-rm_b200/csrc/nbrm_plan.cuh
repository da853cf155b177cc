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

// Strings that come from the reference's process environment / flags.  The
// host JSON-escapes them and appends them to the pool after L__TOTAL.
struct EnvLits {
  Span istio_host;      // ISTIO_HOST or "*"            (notebook_controller.go:548-551)
  Span istio_gateway;   // ISTIO_GATEWAY or default     (:557-560)
  Span cluster_domain;  // CLUSTER_DOMAIN or default    (:537-539)
  Span ctrl_namespace;  // ODH controller namespace     (odh main.go:63-72)
  Span oauth_image;     // --oauth-proxy-image          (odh main.go:80)
  // profile-controller (profile_controller.go:419-504, :127-137)
  Span prof_when_key;   // "request.headers[<userid header>]" (quoted)
  Span prof_prefix;     // userid prefix (unquoted, escaped)
  Span prof_igw, prof_kfp, prof_nbc;  // principals (quoted)
  Span prof_ns_labels;  // members of the Namespace's labels map, rendered by the host
  uint32_t add_fsgroup; // ADD_FSGROUP unset or "true"  (notebook_controller.go:475)
  uint32_t use_istio;   // USE_ISTIO == "true"          (:213)
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
