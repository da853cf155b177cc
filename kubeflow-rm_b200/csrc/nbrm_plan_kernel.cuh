// nbrm_plan_kernel.cuh — the plan kernel template and its launcher.  Each
// generator instantiates it in a translation unit of its own (gen_*.cu) so the
// generators compile in parallel: in one unit the build took 28 minutes.
#pragma once
#include <stdlib.h>
#include <string.h>

#include "nbrm_dev.h"

namespace nbrm {

// One input tuple per thread (document i of every input slot in use).  Writes,
// for each of G::NOBJ output objects, the segment list, its byte length and the
// segment count; plus one status word per tuple.
//
// The documents of a block's tuples are contiguous in each input arena, so the
// block first copies that range of every slot into shared memory with aligned
// 16-byte loads (coalesced, each input byte read from HBM once) and the
// planners then walk shared memory.  Walking global memory directly makes every
// byte access of a warp touch 32 different cache lines, which is what bounded
// the first version of this kernel.  A block whose documents do not fit (a
// Notebook with a huge annotation) walks global memory instead.
extern __shared__ __align__(16) char plan_smem[];

template <class G>
__global__ void plan_kernel(PlanIn in, PlanOut out, uint32_t smem_cap, uint32_t period) {
  const uint32_t b0 = blockIdx.x * blockDim.x;
  const uint32_t b1 = min(b0 + blockDim.x, in.n);
  // Tuples whose kinds repeat with period P (the three objects of one Notebook,
  // one of them ten times the size of the others): threads that run together
  // take tuples of one residue, so they walk documents of one kind and finish
  // together instead of idling behind the largest.  Staged: the block has P
  // warps over 32*P consecutive tuples, warp w takes residue w.  Unstaged: the
  // whole block takes one residue (block b: residue b % P), so the blocks of the
  // small kinds retire at once and the SM fills up with warps of the large one.
  uint32_t i = b0 + threadIdx.x;
  if (period > 1) {
    if (smem_cap)
      i = b0 + (threadIdx.x & 31) * period + (threadIdx.x >> 5);
    else
      i = ((blockIdx.x / period) * blockDim.x + threadIdx.x) * period + blockIdx.x % period;
  }

  // the block's byte range in every slot, widened to 16-byte boundaries
  uint64_t g0[3], a0[3];
  uint32_t soff[3], total = 0;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    g0[k] = a0[k] = 0;
    soff[k] = total;
    if (smem_cap && in.slot[k].data && !(G::SHARED1 && k == 1)) {
      g0[k] = in.slot[k].off[b0];
      const uint64_t g1 = in.slot[k].off[b1];
      a0[k] = g0[k] & ~15ull;
      const uint64_t a1 = (g1 + 15ull) & ~15ull;
      const uint64_t bytes = a1 - a0[k];
      total = bytes > 0xffffffffull - total ? 0xffffffffu : total + (uint32_t)bytes;
    }
  }
  const bool staged = smem_cap != 0 && total <= smem_cap;
  if (staged) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      if (!in.slot[k].data || (G::SHARED1 && k == 1)) continue;
      const uint32_t chunks = ((k < 2 ? soff[k + 1] : total) - soff[k]) >> 4;
      const uint4* src = reinterpret_cast<const uint4*>(in.slot[k].data + a0[k]);
      uint4* dst = reinterpret_cast<uint4*>(plan_smem + soff[k]);
      for (uint32_t c = threadIdx.x; c < chunks; c += blockDim.x) dst[c] = src[c];
    }
  }
  __syncthreads();
  if (i >= in.n) return;
  if (in.perm) i = in.perm[i];  // unstaged, period 1 (launch_thread): tuples in the order of their length

  DocCtx c;
  uint32_t aux_abs = i * AUX_PAD, aux_cap = AUX_PAD;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    if (in.slot[k].data && !(G::SHARED1 && k == 1)) {
      const uint64_t o0 = in.slot[k].off[i], o1 = in.slot[k].off[i + 1];
      c.d[k].p = staged ? plan_smem + soff[k] + (uint32_t)(o0 - a0[k]) : in.slot[k].data + o0;
      c.d[k].n = (uint32_t)(o1 - o0);
      c.base[k] = (uint32_t)o0;
      aux_abs += (uint32_t)o0;
      aux_cap += c.d[k].n;
#ifdef NBRM_TAPE
      if (in.slot[k].tidx) {  // the slot was indexed when it became resident (nbrm_slot_index)
        const uint2 ti = in.slot[k].tidx[i];
        c.d[k].tape = reinterpret_cast<const TapeEnt*>(in.slot[k].tape) + ti.x;
        c.d[k].nt = ti.y & 0xffffu;
        c.d[k].tflags = c.d[k].nt ? (ti.y >> 16) : DOC_WS;
        c.d[k].dlim = TAPE_MAX_DEPTH;
      }
#endif
    } else {
      c.d[k].p = nullptr;
      c.d[k].n = 0;
      c.base[k] = 0;
    }
  }
  c.aux.abs = aux_abs;
  c.aux.p = out.aux + aux_abs;
  c.aux.cap = aux_cap;
  c.aux.used = 0;
  c.env = &in.env;
  c.flags = in.flags;
  c.index = i;
  c.extra = in.extra;
  c.now = in.now;
  c.arg0 = in.arg0;
  c.arg1 = in.arg1;
  c.result = 0;
  c.shared = G::SHARED1 ? in.slot[1].data : nullptr;
  c.shared_off = G::SHARED1 ? in.slot[1].off : nullptr;
  c.shared_n = G::SHARED1 ? in.shared_n : 0;
  c.shared_tape = G::SHARED1 ? in.slot[1].tape : nullptr;
  c.shared_tidx = G::SHARED1 ? (const void*)in.slot[1].tidx : nullptr;
  ObjSink s;
  s.segs = out.segs + (size_t)i * G::NOBJ * out.cap;
  s.nseg = out.nseg + (size_t)i * G::NOBJ;
  s.lens = out.lens + (size_t)i * G::NOBJ;
  s.cap = out.cap;
  s.views.p = out.views;
  s.views.n = in.n * G::NOBJ;
  s.view_j = i * G::NOBJ;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    c.view[k].t.p = (G::SHARED1 && k == 1) ? nullptr : in.slot[k].views;
    c.view[k].t.n = in.n;
    c.view[k].j = i;
  }
  out.status[i] = G::run(c, s);
  out.result[i] = c.result;
}

static const uint32_t PLAN_SMEM = 100 * 1024;  // two blocks per SM

#ifdef NBRM_WARP
// ---------------------------------------------------------------- one tuple per WARP
// Tape mode (nbrm_tape.cuh).  The warp reads each document of its tuple once, 16 bytes
// per lane, coalesced, and leaves the structural index in its share of shared memory;
// then all 32 lanes run the planner together on the same tuple — same branches, same
// addresses, so a load is one transaction and the stores of the plan collapse into
// one — hopping through the index instead of scanning.  Bulk byte work inside the
// planner (span compares) is split across the lanes.  What a warp does depends on its
// own tuple only: a batch of mixed shapes costs the sum of its tuples, not 32 times the
// largest of every 32.
static const uint32_t PLANW_WARPS = 4;
static const uint32_t NBRM_WARP_MAX_TUPLES = 4096;  // measured: profiles/README.md, "warp or thread"

template <class G>
__global__ void __launch_bounds__(PLANW_WARPS * 32) plan_kernel_warp(PlanIn in, PlanOut out,
                                                                       uint32_t tape_cap) {
  extern __shared__ __align__(16) char plan_smem_w[];
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t i = blockIdx.x * PLANW_WARPS + warp;
  if (i >= in.n) return;
  TapeEnt* tape = reinterpret_cast<TapeEnt*>(plan_smem_w) + (size_t)warp * tape_cap;
  TapeScratch* sc =
      reinterpret_cast<TapeScratch*>(plan_smem_w + (size_t)PLANW_WARPS * tape_cap * sizeof(TapeEnt)) + warp;
  DocCtx c;
  uint32_t aux_abs = i * AUX_PAD, aux_cap = AUX_PAD, used = 0;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    if (in.slot[k].data && !(G::SHARED1 && k == 1)) {
      const uint64_t o0 = in.slot[k].off[i], o1 = in.slot[k].off[i + 1];
      c.d[k].p = in.slot[k].data + o0;
      c.d[k].n = (uint32_t)(o1 - o0);
      c.base[k] = (uint32_t)o0;
      aux_abs += (uint32_t)o0;
      aux_cap += c.d[k].n;
      c.d[k].dlim = TAPE_MAX_DEPTH;
      if (in.slot[k].tidx) {  // indexed when it became resident (nbrm_slot_index)
        const uint2 ti = in.slot[k].tidx[i];
        c.d[k].tape = reinterpret_cast<const TapeEnt*>(in.slot[k].tape) + ti.x;
        c.d[k].nt = ti.y & 0xffffu;
        c.d[k].tflags = (c.d[k].nt ? (ti.y >> 16) : DOC_WS) | DOC_COOP;
      } else {
        uint32_t tf = DOC_WS;
        const uint32_t nt = build_tape_warp(in.slot[k].data, o0, c.d[k].n, tape + used, tape_cap - used, sc,
                                            TAPE_MAX_DEPTH, &tf);
        c.d[k].tape = tape + used;
        c.d[k].nt = nt;
        c.d[k].tflags = tf | DOC_COOP;
        used += nt;
      }
    } else {
      c.d[k].p = nullptr;
      c.d[k].n = 0;
      c.base[k] = 0;
    }
  }
  __syncwarp();
  c.aux.abs = aux_abs;
  c.aux.p = out.aux + aux_abs;
  c.aux.cap = aux_cap;
  c.aux.used = 0;
  c.env = &in.env;
  c.flags = in.flags;
  c.index = i;
  c.extra = in.extra;
  c.now = in.now;
  c.arg0 = in.arg0;
  c.arg1 = in.arg1;
  c.result = 0;
  c.shared = G::SHARED1 ? in.slot[1].data : nullptr;
  c.shared_off = G::SHARED1 ? in.slot[1].off : nullptr;
  c.shared_n = G::SHARED1 ? in.shared_n : 0;
  c.shared_tape = G::SHARED1 ? in.slot[1].tape : nullptr;
  c.shared_tidx = G::SHARED1 ? (const void*)in.slot[1].tidx : nullptr;
  ObjSink s;
  s.segs = out.segs + (size_t)i * G::NOBJ * out.cap;
  s.nseg = out.nseg + (size_t)i * G::NOBJ;
  s.lens = out.lens + (size_t)i * G::NOBJ;
  s.cap = out.cap;
  s.views.p = out.views;
  s.views.n = in.n * G::NOBJ;
  s.view_j = i * G::NOBJ;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    c.view[k].t.p = (G::SHARED1 && k == 1) ? nullptr : in.slot[k].views;
    c.view[k].t.n = in.n;
    c.view[k].j = i;
  }
  const uint32_t st = G::run(c, s);
  if ((threadIdx.x & 31u) == 0) {
    out.status[i] = st;
    out.result[i] = c.result;
  }
}

// Tape entries per warp: JSON of API objects has one structural character per 8 to 11
// bytes; tuples whose kinds repeat with a period (one Notebook's StatefulSet, Service and
// VirtualService) are far from their average size, hence the factor.  A tuple that does
// not fit keeps the documents that did and walks the others as text.
static uint32_t tape_cap_for(uint64_t in_bytes, uint32_t n, uint32_t period) {
  const uint64_t avg = in_bytes / n + 64;
  uint64_t cap = (period > 1 ? avg * 5 / 2 : avg * 3 / 2) / 8;
  if (cap < 192) cap = 192;
  if (cap > 1536) cap = 1536;
  return (uint32_t)((cap + 31) & ~31ull);
}

template <class G>
static void launch_warp(const PlanIn& in, const PlanOut& out, uint64_t in_bytes, cudaStream_t st) {
  if (!in.n) return;
  const uint32_t cap = tape_cap_for(in_bytes, in.n, in.period);
  const size_t smem = (size_t)PLANW_WARPS * (cap * sizeof(TapeEnt) + sizeof(TapeScratch));
  cudaFuncSetAttribute(plan_kernel_warp<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const uint32_t blocks = (in.n + PLANW_WARPS - 1) / PLANW_WARPS;
  plan_kernel_warp<G><<<blocks, PLANW_WARPS * 32, smem, st>>>(in, out, cap);
}
#endif  // NBRM_WARP

template <class G>
static void launch_thread(const PlanIn& in, const PlanOut& out, uint64_t in_bytes, cudaStream_t st) {
  if (!in.n) return;
  // opt in to 100 KB of dynamic shared memory; the attribute belongs to the current
  // device's context, so it is set on every launch (a process may drive several GPUs)
  cudaFuncSetAttribute(plan_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PLAN_SMEM);
  // Staging pays while a block's tuples (one per thread) fit: 128 or 64 tuples in
  // 100 KB.  Larger tuples would leave one warp per block and two blocks per SM,
  // far too few threads to hide the walk's latency; those run unstaged, 128
  // threads per block with no shared memory, on the L1/L2 caches.
#ifdef NBRM_NO_STAGING
  const uint64_t avg = PLAN_SMEM;  // this unit's kernel always walks global memory
#else
  const uint64_t avg = in.perm ? PLAN_SMEM : in_bytes / in.n + 48;  // a permuted batch is not contiguous per block
#endif
  if (in.period > 1 && in.period <= 8) {  // threads that run together take one kind (see plan_kernel)
    if ((uint64_t)32 * in.period * avg <= PLAN_SMEM) {
      const uint32_t threads = 32 * in.period;
      plan_kernel<G><<<(in.n + threads - 1) / threads, threads, PLAN_SMEM, st>>>(in, out, PLAN_SMEM,
                                                                                 in.period);
    } else {
      const uint32_t per_kind = (in.n + in.period - 1) / in.period;
      const uint32_t blocks = ((per_kind + 127) / 128) * in.period;
      plan_kernel<G><<<blocks, 128, 0, st>>>(in, out, 0, in.period);
    }
    return;
  }
  uint32_t threads = 128;
  while (threads > 64 && (uint64_t)threads * avg > PLAN_SMEM) threads >>= 1;
  if ((uint64_t)threads * avg > PLAN_SMEM) {
    uint32_t blocks = (in.n + 127) / 128;
    plan_kernel<G><<<blocks, 128, 0, st>>>(in, out, 0, 1);
    return;
  }
  uint32_t blocks = (in.n + threads - 1) / threads;
  plan_kernel<G><<<blocks, threads, PLAN_SMEM, st>>>(in, out, PLAN_SMEM, 1);
}

// A thread per tuple does 32 tuples' worth of work per instruction once the GPU is full of
// them, a warp per tuple finishes one tuple 30 times sooner: small batches (a latency
// question) go to the warp kernel, large ones (throughput) to the thread kernel.
// NBRM_PLAN_MODE=warp|thread overrides the choice (experiments, tests).
template <class G>
static void launch(const PlanIn& in, const PlanOut& out, uint64_t in_bytes, cudaStream_t st) {
#ifdef NBRM_WARP
  static const int mode = [] {
    const char* e = getenv("NBRM_PLAN_MODE");
    return !e ? 0 : !strcmp(e, "warp") ? 1 : !strcmp(e, "thread") ? 2 : 0;
  }();
  if (mode == 1 || (mode == 0 && in.n <= NBRM_WARP_MAX_TUPLES)) return launch_warp<G>(in, out, in_bytes, st);
#endif
  launch_thread<G>(in, out, in_bytes, st);
}

// Defines the two entry points nbrm_plan_kernels.cu dispatches to.
#define NBRM_DEFINE_GENERATOR(NAME, G)                                                        \
  void launch_plan_##NAME(const PlanIn& in, const PlanOut& out, uint64_t in_bytes,           \
                          cudaStream_t st) {                                                  \
    launch<G>(in, out, in_bytes, st);                                                         \
  }                                                                                           \
  void gen_shape_##NAME(uint32_t* nobj, uint32_t* cap, bool* shared1) {                       \
    *nobj = G::NOBJ;                                                                          \
    *cap = G::CAP;                                                                            \
    *shared1 = G::SHARED1;                                                                    \
  }

}  // namespace nbrm
