// nbrm_plan_kernels.cu — the plan kernel, instantiated once per generator.
#include "nbrm_dev.h"
#include "nbrm_nbc.cuh"
#include "nbrm_odh.cuh"
#include "nbrm_cmp.cuh"

namespace nbrm {

// One input tuple per thread (document i of every input slot in use).  Writes,
// for each of G::NOBJ output objects, the segment list, its byte length and the
// segment count; plus one status word per tuple.
template <class G>
__global__ void __launch_bounds__(PLAN_THREADS) plan_kernel(PlanIn in, PlanOut out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= in.n) return;
  DocCtx c;
  uint32_t aux_abs = i * AUX_PAD, aux_cap = AUX_PAD;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    if (in.slot[k].data) {
      const uint64_t o0 = in.slot[k].off[i], o1 = in.slot[k].off[i + 1];
      c.d[k].p = in.slot[k].data + o0;
      c.d[k].n = (uint32_t)(o1 - o0);
      c.base[k] = (uint32_t)o0;
      aux_abs += (uint32_t)o0;
      aux_cap += c.d[k].n;
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
  ObjSink s;
  s.segs = out.segs + (size_t)i * G::NOBJ * out.cap;
  s.nseg = out.nseg + (size_t)i * G::NOBJ;
  s.lens = out.lens + (size_t)i * G::NOBJ;
  s.cap = out.cap;
  out.status[i] = G::run(c, s);
}

template <class G>
static void launch(const PlanIn& in, const PlanOut& out, cudaStream_t st) {
  if (!in.n) return;
  uint32_t blocks = (in.n + PLAN_THREADS - 1) / PLAN_THREADS;
  plan_kernel<G><<<blocks, PLAN_THREADS, 0, st>>>(in, out);
}

void launch_plan(int gen, const PlanIn& in, const PlanOut& out, cudaStream_t st) {
  switch (gen) {
    case GEN_ID_NBC: launch<GenNbc>(in, out, st); break;
    case GEN_ID_ODH: launch<GenOdh>(in, out, st); break;
    case GEN_ID_COPY: launch<GenCopyFields>(in, out, st); break;
  }
}

void gen_shape(int gen, uint32_t* nobj, uint32_t* cap) {
  switch (gen) {
    case GEN_ID_NBC: *nobj = GenNbc::NOBJ; *cap = GenNbc::CAP; break;
    case GEN_ID_ODH: *nobj = GenOdh::NOBJ; *cap = GenOdh::CAP; break;
    case GEN_ID_COPY: *nobj = GenCopyFields::NOBJ; *cap = GenCopyFields::CAP; break;
    default: *nobj = 0; *cap = 0;
  }
}

}  // namespace nbrm
