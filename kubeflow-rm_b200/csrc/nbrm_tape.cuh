// nbrm_tape.cuh — the structural index ("tape") of one JSON document.
//
// A thread that walks a document alone spends its time finding where values end: every
// byte of a skipped container is looked at, and the 32 lanes of a warp do so in 32
// different cache lines.  In tape mode (NBRM_TAPE) a WARP owns one input tuple.  It first
// reads each document once, 16 bytes per lane with coalesced loads, and writes down every
// structural character that is not inside a string — { } [ ] : , — with its position and
// nesting depth, plus two links per entry.  The planners then hop from member to member
// through the links instead of scanning: ObjIter / ArrIter / Cursor in nbrm_json.cuh use
// the tape when the document has one and fall back to the text walk when it has not.
//
//   w    = pos (21 bits) | kind (4 bits) << 21 | depth (7 bits) << 25
//          kind = (c & 7) | ((c & 0x20) >> 2)  of the character c:
//          '{' 11  '[' 3  '}' 13  ']' 5  ':' 10  ',' 12   ((kind & 7): 3 open, 5 close, 2 colon, 4 comma)
//          depth = nesting depth of the container the character belongs to (root = 1;
//          a bracket belongs to the container it opens or closes)
//   link = low 16 bits:  index of the next ',' or closer at the same depth, i.e. where the
//                        value that starts after this entry ends          (0xffff: none)
//          high 16 bits: for an opener, the index of its closer           (0xffff: none)
//
// Limits of the tape (a document beyond them simply has none and is walked as text):
// 2 MiB per document, nesting depth 63, and the entries of one tuple must fit the warp's
// share of shared memory.
//
// This header holds the format, the scalar reference builder (host: the CPU build of the
// planners used by the CPU tests, and the check of the device builder) and the device builder.
#pragma once
#include <stdint.h>

namespace nbrm {

static const uint32_t TAPE_POS_MASK = (1u << 21) - 1;
static const uint32_t TAPE_MAX_DOC = 1u << 21;
static const uint32_t TAPE_MAX_DEPTH = 63;
static const uint32_t TAPE_NOLINK = 0xffffu;
static const uint32_t DOC_WS = 1u;    // Doc::tflags: whitespace (or control bytes) between tokens
static const uint32_t DOC_COOP = 2u;  // all 32 lanes of the warp walk this document together (plan_kernel_warp)

NB_HD uint32_t tpos(uint32_t w) { return w & TAPE_POS_MASK; }
NB_HD uint32_t tkind(uint32_t w) { return (w >> 21) & 7u; }
NB_HD uint32_t tdepth(uint32_t w) { return w >> 25; }
NB_HD bool t_open(uint32_t w) { return tkind(w) == 3u; }
NB_HD bool t_close(uint32_t w) { return tkind(w) == 5u; }
NB_HD bool t_colon(uint32_t w) { return tkind(w) == 2u; }
NB_HD bool t_comma(uint32_t w) { return tkind(w) == 4u; }
NB_HD uint32_t t_sib(const TapeEnt& e) { return e.link & 0xffffu; }
NB_HD uint32_t t_match(const TapeEnt& e) { return e.link >> 16; }
NB_HD uint32_t tape_word(uint32_t pos, unsigned char c, uint32_t depth) {
  return pos | ((((uint32_t)c & 7u) | (((uint32_t)c & 0x20u) >> 2)) << 21) | (depth << 25);
}

// First tape index whose position is >= pos (nt when there is none).  `guess` is tried
// first: walks move forward, so the caller's last index is usually the answer.
NB_HD uint32_t tape_lower_bound(const Doc& d, uint32_t pos, uint32_t guess = TNONE) {
  if (guess < d.nt && tpos(d.tape[guess].w) >= pos && (guess == 0 || tpos(d.tape[guess - 1].w) < pos))
    return guess;
  uint32_t lo = 0, hi = d.nt;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (tpos(d.tape[mid].w) < pos) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// ---------------------------------------------------------------- scalar builder
// One pass with a string / escape state, then the links from a stack per depth.  Returns
// the number of entries, 0 when the document cannot have a tape.  The device builder
// below must produce exactly these entries (tests/test_gpu_tape.py).
inline uint32_t build_tape_scalar(const char* p, uint32_t n, TapeEnt* out, uint32_t cap, uint32_t* tflags) {
  *tflags = DOC_WS;
  if (n >= TAPE_MAX_DOC) return 0;
  uint32_t nt = 0, depth = 0, flags = 0;
  bool in_str = false, esc = false;
  for (uint32_t i = 0; i < n; ++i) {
    const unsigned char c = (unsigned char)p[i];
    if (in_str) {
      if (esc) esc = false;
      else if (c == '\\') esc = true;
      else if (c == '"') in_str = false;
      continue;
    }
    if (c == '"') {
      in_str = true;
      continue;
    }
    if (c <= 0x20) flags |= DOC_WS;
    const bool open = c == '{' || c == '[', close = c == '}' || c == ']';
    if (!open && !close && c != ':' && c != ',') continue;
    if (open) ++depth;
    if (depth > TAPE_MAX_DEPTH || nt >= cap || nt >= TAPE_NOLINK) return 0;
    out[nt].w = tape_word(i, c, depth);
    out[nt].link = 0xffffffffu;
    ++nt;
    if (close) {
      if (!depth) return 0;  // more closers than openers: not a document the tape can describe
      --depth;
    }
  }
  // links, back to front: the nearest later ',' / closer and the nearest later closer per depth
  uint32_t last_sep[TAPE_MAX_DEPTH + 1], last_close[TAPE_MAX_DEPTH + 1];
  for (uint32_t k = 0; k <= TAPE_MAX_DEPTH; ++k) last_sep[k] = last_close[k] = TAPE_NOLINK;
  for (uint32_t k = nt; k-- > 0;) {
    const uint32_t w = out[k].w, dp = tdepth(w);
    const uint32_t sib = t_close(w) ? TAPE_NOLINK : last_sep[dp];
    const uint32_t mt = t_open(w) ? last_close[dp] : TAPE_NOLINK;
    out[k].link = sib | (mt << 16);
    if (t_close(w) || t_comma(w)) last_sep[dp] = k;
    if (t_close(w)) last_close[dp] = k;
  }
  *tflags = flags;
  return nt;
}

#if defined(__CUDACC__)
// ---------------------------------------------------------------- warp builder
// bit 7 of every byte of x that equals the byte replicated in c4 (exact; the other bits are
// garbage and are masked by the caller)
__device__ __forceinline__ uint32_t swar_eq(uint32_t x, uint32_t c4) {
  const uint32_t a = x ^ c4;
  return ~(((a & 0x7f7f7f7fu) + 0x7f7f7f7fu) | a);
}

// Per-warp scratch of the link pass: the nearest later separator / closer per depth.
struct TapeScratch {
  uint16_t last_sep[TAPE_MAX_DEPTH + 1];
  uint16_t last_close[TAPE_MAX_DEPTH + 1];
};

// Builds the tape of the document arena[o0, o0+n) into `tape` (shared memory, room for
// `cap` entries).  Called by all 32 lanes of a warp; returns the entry count (0: no tape)
// to every lane.  dlim: lanes whose 16 bytes lie entirely deeper than this write nothing
// (entries of depth <= dlim are always complete; TAPE_MAX_DEPTH keeps everything).
static __device__ __noinline__ uint32_t build_tape_warp(const char* __restrict__ arena, uint64_t o0, uint32_t n,
                                                 TapeEnt* tape, uint32_t cap, TapeScratch* sc,
                                                 uint32_t dlim, uint32_t* tflags) {
  const uint32_t FULL = 0xffffffffu, K80 = 0x80808080u;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t lt = (1u << lane) - 1u;
  *tflags = DOC_WS;
  if (n == 0 || n >= TAPE_MAX_DOC) return 0;
  const uint64_t end = o0 + n;
  const uint64_t a0 = o0 & ~15ull;
  const uint32_t chunks = (uint32_t)((end - a0 + 511) >> 9);
  uint32_t depth = 0, nt = 0, instr = 0, esc_carry = 0, wsany = 0;
  bool fail = false;
  uint64_t la = a0 + lane * 16u;
  uint4 v = make_uint4(0x21212121u, 0x21212121u, 0x21212121u, 0x21212121u);
  if (la < end) v = *reinterpret_cast<const uint4*>(arena + la);
  for (uint32_t j = 0; j < chunks; ++j) {
    uint32_t x[4] = {v.x, v.y, v.z, v.w};
    const uint64_t cur = la;
    // the next chunk's bytes are on their way while this one is classified
    la += 512;
    v = make_uint4(0x21212121u, 0x21212121u, 0x21212121u, 0x21212121u);
    if (j + 1 < chunks && la < end) v = *reinterpret_cast<const uint4*>(arena + la);
    if (cur < o0 || cur + 16 > end) {  // bytes of the neighbouring documents count as '!'
      const uint32_t lo = cur < o0 ? (uint32_t)(o0 - cur) : 0u;
      const uint32_t hi = cur >= end ? 0u : (cur + 16 > end ? (uint32_t)(end - cur) : 16u);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        uint32_t m = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b)
          if ((uint32_t)(4 * k + b) >= lo && (uint32_t)(4 * k + b) < hi) m |= 0xffu << (8 * b);
        x[k] = (x[k] & m) | (0x21212121u & ~m);
      }
    }
    uint32_t q[4], S[4], O[4], C[4], bsany = 0, W[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t y = x[k] | 0x20202020u;
      q[k] = swar_eq(x[k], 0x22222222u) & K80;
      bsany |= swar_eq(x[k], 0x5c5c5c5cu);
      O[k] = swar_eq(y, 0x7b7b7b7bu);
      C[k] = swar_eq(y, 0x7d7d7d7du);
      S[k] = O[k] | C[k] | swar_eq(x[k], 0x3a3a3a3au) | swar_eq(x[k], 0x2c2c2c2cu);
      W[k] = ~(((x[k] & 0x7f7f7f7fu) + 0x5f5f5f5fu) | x[k]);  // bytes <= 0x20
    }
    // escapes (rare): a quote after an odd run of backslashes does not end the string
    if (__ballot_sync(FULL, (bsany & K80) != 0) | esc_carry) {
      uint32_t bs[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) bs[k] = swar_eq(x[k], 0x5c5c5c5cu) & K80;
      uint32_t t = 0;  // backslashes at the end of this lane's bytes
#pragma unroll
      for (int i = 15; i >= 0; --i) {
        if (!((bs[i >> 2] >> (8 * (i & 3) + 7)) & 1u)) break;
        ++t;
      }
      const uint32_t fullm = __ballot_sync(FULL, t == 16), oddm = __ballot_sync(FULL, t & 1u);
      const uint32_t below = ~fullm & lt;  // lanes before this one that hold something else too
      uint32_t e = below ? ((oddm >> (31 - __clz(below))) & 1u) : esc_carry;
      uint32_t E[4] = {0, 0, 0, 0};
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const uint32_t bit = 1u << (8 * (i & 3) + 7);
        if (e) {
          E[i >> 2] |= bit;
          e = 0;
        } else if (bs[i >> 2] & bit) {
          e = 1;
        }
      }
      esc_carry = __shfl_sync(FULL, e, 31);
#pragma unroll
      for (int k = 0; k < 4; ++k) q[k] &= ~E[k];
    }
    // inside-a-string flags: prefix parity of the quotes, first inside the lane ...
    uint32_t ins[4], inc = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      ins[k] = (q[k] * 0x01010101u) ^ inc;
      inc = (uint32_t)((int32_t)ins[k] >> 31);
    }
    // ... then across the lanes and the chunks before
    const uint32_t parm = __ballot_sync(FULL, inc & 1u);
    const uint32_t in_abs = ((__popc(parm & lt) & 1u) ^ instr) ? FULL : 0u;
    instr ^= __popc(parm) & 1u;
    uint32_t wl = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t out = ~(ins[k] ^ in_abs) & K80;  // bytes outside strings
      S[k] &= out;
      O[k] &= out;
      C[k] &= out;
      wl |= W[k] & out;
    }
    wsany |= wl;
    // depth before this lane's bytes and rank of its first entry: one scan of (count, net)
    const uint32_t po = __popc(O[0] | (O[1] >> 1) | (O[2] >> 2) | (O[3] >> 3));
    const uint32_t pc = __popc(C[0] | (C[1] >> 1) | (C[2] >> 2) | (C[3] >> 3));
    uint32_t cnt = __popc(S[0] | (S[1] >> 1) | (S[2] >> 2) | (S[3] >> 3));
    uint32_t net = po - pc + 16u;  // biased: 0..32
    uint32_t inc_net = net;
#pragma unroll
    for (int dd = 1; dd < 32; dd <<= 1) {
      const uint32_t t = __shfl_up_sync(FULL, inc_net, dd);
      if (lane >= (uint32_t)dd) inc_net += t;
    }
    int dep = (int)depth + (int)(inc_net - net) - 16 * (int)lane;  // depth before this lane
    depth = (uint32_t)((int)depth + (int)__shfl_sync(FULL, inc_net, 31) - 16 * 32);
    if (dep - (int)pc > (int)dlim) cnt = 0;  // everything here is deeper than anyone will look
    if (__ballot_sync(FULL, cnt != 0)) {
      uint32_t inc_cnt = cnt;
#pragma unroll
      for (int dd = 1; dd < 32; dd <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL, inc_cnt, dd);
        if (lane >= (uint32_t)dd) inc_cnt += t;
      }
      const uint32_t total = __shfl_sync(FULL, inc_cnt, 31);
      if (nt + total > cap || nt + total >= TAPE_NOLINK) fail = true;
      if (!fail && cnt) {
        uint32_t r = nt + inc_cnt - cnt;
        const uint32_t pos0 = (uint32_t)(cur - o0);  // wraps for the lane that starts before the document: its
                                                      // entries all lie inside, so pos0 + offset is right again
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          uint32_t s = S[k];
          while (s) {
            const uint32_t bit = __ffs(s) - 1;  // 7, 15, 23, 31
            s &= s - 1;
            const uint32_t c = (x[k] >> (bit - 7)) & 0xffu;
            const uint32_t kind = (c & 7u) | ((c & 0x20u) >> 2);
            dep += (kind & 7u) == 3u;
            if (dep < 0 || dep > (int)TAPE_MAX_DEPTH) fail = true;
            TapeEnt e;
            e.w = (pos0 + 4 * k + (bit >> 3)) | (kind << 21) | ((uint32_t)(dep & 127) << 25);
            e.link = 0xffffffffu;
            tape[r++] = e;
            dep -= (kind & 7u) == 5u;
            if (dep < 0) fail = true;  // a closer without its opener
          }
        }
      }
      nt += total;
    }
    if (__ballot_sync(FULL, fail)) return 0;
  }
  if (nt == 0) {
    *tflags = (__ballot_sync(FULL, (wsany & K80) != 0)) ? DOC_WS : 0u;
    return 0;
  }
  // links, back to front, 32 entries at a time
  for (uint32_t k = lane; k <= TAPE_MAX_DEPTH; k += 32) sc->last_sep[k] = sc->last_close[k] = (uint16_t)TAPE_NOLINK;
  __syncwarp();
  const uint32_t gt = ~lt & ~(1u << lane);
  for (uint32_t base = (nt - 1) & ~31u;; base -= 32) {
    const uint32_t idx = base + lane;
    const bool valid = idx < nt;
    const uint32_t w = valid ? tape[idx].w : 0u;
    const uint32_t dp = tdepth(w);
    const bool cl = valid && t_close(w), tg = cl || (valid && t_comma(w));
    const uint32_t peers = __match_any_sync(FULL, valid ? dp : 128u + lane);
    const uint32_t tgm = __ballot_sync(FULL, tg), clm = __ballot_sync(FULL, cl);
    if (valid) {
      const uint32_t c1 = peers & tgm & gt, c2 = peers & clm & gt;
      const uint32_t sib = cl ? TAPE_NOLINK : (c1 ? base + __ffs(c1) - 1 : sc->last_sep[dp]);
      const uint32_t mt = t_open(w) ? (c2 ? base + __ffs(c2) - 1 : sc->last_close[dp]) : TAPE_NOLINK;
      tape[idx].link = sib | (mt << 16);
    }
    __syncwarp();
    if (tg && !(peers & tgm & lt)) sc->last_sep[dp] = (uint16_t)idx;
    if (cl && !(peers & clm & lt)) sc->last_close[dp] = (uint16_t)idx;
    __syncwarp();
    if (base == 0) break;
  }
  *tflags = (__ballot_sync(FULL, (wsany & K80) != 0)) ? DOC_WS : 0u;
  return nt;
}
#endif  // __CUDACC__

}  // namespace nbrm
