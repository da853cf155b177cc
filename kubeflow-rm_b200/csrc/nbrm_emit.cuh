// nbrm_emit.cuh — the per-lane work of the emit kernel (emit_warp, nbrm_kernels.cu).
//
// emit_warp gives an output object to a warp and walks its bytes in rounds of 32 aligned
// 16-byte blocks of the OUTPUT address space (512 bytes), lane l on block l.  A round that
// lies inside one segment (the inside of a copied PodSpec, template text) is copied source to
// output with one shift that is the same for the whole warp.  Otherwise a lane finds the
// segment its first byte lies in (a binary search over the few segments that start inside the
// round; their prefix sums are in shared memory) and gathers its 16 bytes from however many
// segments they span: every piece is read through the aligned 16-byte blocks of its source,
// moved with ONE shift from its alignment in the source to its place in the block, and laid
// over the block in registers.  A warp writes 512 contiguous bytes per store instruction and
// no lane waits for another lane's loads.  Only the first and last block of an object are
// shared with its neighbours in the packed output and are stored in smaller pieces.
//
// The routines are __host__ __device__ so that the CPU test build of the planner headers can
// run them block by block against a plain concatenation.
#pragma once
#include <stdint.h>
#include <string.h>

#include "nbrm_plan.cuh"

namespace nbrm {
namespace emitw {

// 16 bytes, little endian: byte k of the block is bits 8k.. of lo (k < 8) or 8(k-8).. of hi
struct B16 {
  uint64_t lo, hi;
};

NB_HD B16 load16(const char* aligned16) {
  B16 r;
#ifdef __CUDA_ARCH__
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(aligned16));
  r.lo = (uint64_t)v.x | ((uint64_t)v.y << 32);
  r.hi = (uint64_t)v.z | ((uint64_t)v.w << 32);
#else
  memcpy(&r, aligned16, 16);
#endif
  return r;
}

NB_HD void store16(char* dst, B16 v) {
#ifdef __CUDA_ARCH__
  uint4 w;
  w.x = (uint32_t)v.lo;
  w.y = (uint32_t)(v.lo >> 32);
  w.z = (uint32_t)v.hi;
  w.w = (uint32_t)(v.hi >> 32);
  *reinterpret_cast<uint4*>(dst) = w;
#else
  memcpy(dst, &v, 16);
#endif
}

// bytes k .. k+15 of the 32 bytes (a, b); k in 0..15
NB_HD B16 shr_bytes(B16 a, B16 b, uint32_t k) {
  uint64_t x0 = a.lo, x1 = a.hi, x2 = b.lo;
  if (k & 8u) {
    x0 = a.hi;
    x1 = b.lo;
    x2 = b.hi;
  }
  const uint32_t sh = (k & 7u) * 8u;
  B16 r;
  r.lo = (x0 >> sh) | ((x1 << 1) << (63u - sh));  // (x << 1) << (63 - sh): a shift by 64 that gives 0
  r.hi = (x1 >> sh) | ((x2 << 1) << (63u - sh));
  return r;
}

// Largest g in [lo, hi] with P[g] <= q (P ascending, P[lo] <= q).
NB_HD uint32_t find_seg(const uint32_t* P, uint32_t lo, uint32_t hi, uint32_t q) {
  while (lo < hi) {
    const uint32_t mid = (lo + hi + 1u) >> 1;
    if (P[mid] <= q) lo = mid;
    else hi = mid - 1u;
  }
  return lo;
}

// bytes [f0, f1) of the aligned 16-byte block at dst take their value from v; the other bytes
// of the block belong to someone else and are not touched
NB_HD void store_part(char* dst, B16 v, uint32_t f0, uint32_t f1) {
  if (f0 == 0 && f1 == 16u) {
    store16(dst, v);
    return;
  }
  for (uint32_t w = 0; w < 4u; ++w) {
    const uint32_t b0 = 4u * w;
    if (b0 + 4u <= f0 || b0 >= f1) continue;
    const uint32_t word = (uint32_t)((w & 2u ? v.hi : v.lo) >> (32u * (w & 1u)));
    if (b0 >= f0 && b0 + 4u <= f1) {
#ifdef __CUDA_ARCH__
      *reinterpret_cast<uint32_t*>(dst + b0) = word;
#else
      memcpy(dst + b0, &word, 4);
#endif
    } else {
      for (uint32_t k = 0; k < 4u; ++k)
        if (b0 + k >= f0 && b0 + k < f1) dst[b0 + k] = (char)(word >> (8u * k));
    }
  }
}

// Block i of a run of blocks that lie inside one segment: src is where the run's first byte
// is in the source.  The shift is the same for every block of the run.
NB_HD void copy_pure_block(const char* src, char* dst, uint32_t i) {
  const uint32_t k = (uint32_t)((uintptr_t)src & 15u);
  const char* a = src - k + 16ull * i;
  B16 v = load16(a);
  if (k) v = shr_bytes(v, load16(a + 16), k);
  store16(dst + 16ull * i, v);
}

// ---- a piece in 32-bit words ---------------------------------------------------------------
// X: 12 words — four of anything, the aligned source block that holds the piece's first byte,
// the block after it (needed only when the piece reaches into it).  The piece starts at byte k
// of the first block and goes to byte f of the output block: output word w is bytes
// [4w + e, 4w + e + 4) of X with e = 16 + k - f.  The words come through memory (shared
// memory on the device) because which five of them are needed depends on e.
NB_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t bits) {
#ifdef __CUDA_ARCH__
  return __funnelshift_r(lo, hi, bits);
#else
  return (uint32_t)((((uint64_t)hi << 32) | lo) >> (bits & 31u));
#endif
}

NB_HD void shift_words(const uint32_t* X, uint32_t e, uint32_t o[4]) {
  const uint32_t* x = X + (e >> 2);
  const uint32_t bits = (e & 3u) * 8u;
  const uint32_t x0 = x[0], x1 = x[1], x2 = x[2], x3 = x[3], x4 = x[4];
  o[0] = funnel_r(x0, x1, bits);
  o[1] = funnel_r(x1, x2, bits);
  o[2] = funnel_r(x2, x3, bits);
  o[3] = funnel_r(x3, x4, bits);
}

// word w of "all bytes below position x (0..16)": the table T[x][w] the kernel keeps in shared memory
NB_HD uint32_t below_word(uint32_t x, uint32_t w) {
  if (x >= 4u * w + 4u) return ~0u;
  if (x <= 4u * w) return 0u;
  return (1u << (8u * (x - 4u * w))) - 1u;
}

// One block gathered by one lane.  P[0..m]: prefix sums of the lengths of segments S[0..m)
// (P[0] = 0).  The block at dst gets bytes [f, f + (q_hi - q)) from positions [q, q_hi) of the
// concatenation; g is the segment position q lies in.  X: the lane's 12 words of scratch;
// T: the table of below_word (T[4 x + w]).  Every piece is laid over everything from f upwards,
// so what it carries beyond its own bytes is overwritten by the next piece.
NB_HD void emit_block(const uint32_t* P, const Seg* S, uint32_t g, const char* const* bases, uint32_t q,
                      uint32_t q_hi, uint32_t f, char* dst, uint32_t* X, const uint32_t* T) {
  const uint32_t f0 = f;
  uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
  uint32_t p0 = P[g];
  while (q < q_hi) {
    const Seg sg = S[g];
    const uint32_t p1 = P[g + 1];
    uint32_t nb = p1 - q;
    if (nb > q_hi - q) nb = q_hi - q;
    const char* src = bases[sg.len_src >> 28] + sg.off + (q - p0);
    const uint32_t k = (uint32_t)((uintptr_t)src & 15u);
    const char* blk = src - k;
#ifdef __CUDA_ARCH__
    *reinterpret_cast<uint4*>(X + 4) = __ldg(reinterpret_cast<const uint4*>(blk));
    if (k + nb > 16u) *reinterpret_cast<uint4*>(X + 8) = __ldg(reinterpret_cast<const uint4*>(blk + 16));
#else
    memcpy(X + 4, blk, 16);
    if (k + nb > 16u) memcpy(X + 8, blk + 16, 16);
#endif
    uint32_t o[4];
    shift_words(X, 16u + k - f, o);
    const uint32_t* m = T + 4u * f;
    a0 = (a0 & m[0]) | (o[0] & ~m[0]);
    a1 = (a1 & m[1]) | (o[1] & ~m[1]);
    a2 = (a2 & m[2]) | (o[2] & ~m[2]);
    a3 = (a3 & m[3]) | (o[3] & ~m[3]);
    q += nb;
    f += nb;
    p0 = p1;
    ++g;
  }
  B16 v;
  v.lo = (uint64_t)a0 | ((uint64_t)a1 << 32);
  v.hi = (uint64_t)a2 | ((uint64_t)a3 << 32);
  store_part(dst, v, f0, f);
}

}  // namespace emitw
}  // namespace nbrm
