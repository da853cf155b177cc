// nbrm_nbc.cu — plan kernel for the notebook-controller generators.
#include "nbrm_dev.h"
#include "nbrm_nbc.cuh"

namespace nbrm {

// One Notebook per thread.  Writes, for each of the three output objects, the
// segment list, its length in bytes and the segment count; plus one status
// word per Notebook.
__global__ void __launch_bounds__(PLAN_THREADS) nbc_generate_plan(DevDocs nb, EnvLits env,
                                                                  uint32_t flags, PlanOut out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nb.n) return;
  const uint64_t o0 = nb.off[i], o1 = nb.off[i + 1];
  Doc d;
  d.p = nb.data + o0;
  d.n = (uint32_t)(o1 - o0);

  NbView v = view_notebook(d);
  uint32_t st = v.err;
  if (!st && (!is_object(d, v.podspec) || !v.c0.ok())) st = E_NO_CONTAINERS;

  Aux aux;
  aux.abs = (uint32_t)o0 + i * AUX_PAD;
  aux.p = out.aux + aux.abs;
  aux.cap = d.n + AUX_PAD;
  aux.used = 0;

  Planner P;
  P.base[0] = (uint32_t)o0;
  P.base[1] = P.base[2] = 0;
#pragma unroll 1
  for (uint32_t k = 0; k < NBC_OUT_PER_DOC; ++k) {
    const uint32_t oi = i * NBC_OUT_PER_DOC + k;
    P.begin(out.segs + (size_t)oi * out.cap, out.cap);
    P.err = E_OK;
    if (!st && (flags & (1u << k))) {
      if (k == 0) plan_statefulset(P, d, v, env, flags);
      else if (k == 1) plan_service(P, d, v, flags);
      else plan_virtualservice(P, d, v, env, flags, aux);
      if (P.err) st = P.err;
    }
    const bool ok = !st && (flags & (1u << k));
    out.nseg[oi] = ok ? P.n : 0;
    out.lens[oi] = ok ? P.total : 0;
  }
  if (st) {  // one failing object voids all three
    for (uint32_t k = 0; k < NBC_OUT_PER_DOC; ++k) {
      out.nseg[i * NBC_OUT_PER_DOC + k] = 0;
      out.lens[i * NBC_OUT_PER_DOC + k] = 0;
    }
  }
  out.status[i] = st;
}

void launch_nbc_generate_plan(const DevDocs& nb, const EnvLits& env, uint32_t flags,
                              const PlanOut& out, cudaStream_t st) {
  if (!nb.n) return;
  uint32_t blocks = (nb.n + PLAN_THREADS - 1) / PLAN_THREADS;
  nbc_generate_plan<<<blocks, PLAN_THREADS, 0, st>>>(nb, env, flags, out);
}

}  // namespace nbrm
