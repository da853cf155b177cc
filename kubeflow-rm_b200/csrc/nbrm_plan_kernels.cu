// nbrm_plan_kernels.cu — dispatch from a generator id to its plan kernel.  The
// kernels themselves are instantiated in gen_*.cu, one unit per generator.
#include "nbrm_dev.h"

namespace nbrm {

void launch_plan_nbc(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_nbc(uint32_t*, uint32_t*, bool*);
void launch_plan_odh(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_odh(uint32_t*, uint32_t*, bool*);
void launch_plan_copy(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_copy(uint32_t*, uint32_t*, bool*);
void launch_plan_status(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_status(uint32_t*, uint32_t*, bool*);
void launch_plan_cull(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_cull(uint32_t*, uint32_t*, bool*);
void launch_plan_mutate(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_mutate(uint32_t*, uint32_t*, bool*);
void launch_plan_guard(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_guard(uint32_t*, uint32_t*, bool*);
void launch_plan_poddefault(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_poddefault(uint32_t*, uint32_t*, bool*);
void launch_plan_profile(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_profile(uint32_t*, uint32_t*, bool*);
void launch_plan_odhdecide(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_odhdecide(uint32_t*, uint32_t*, bool*);

void launch_plan_diff(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_diff(uint32_t*, uint32_t*, bool*);
void launch_plan_reviewopen(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_reviewopen(uint32_t*, uint32_t*, bool*);
void launch_plan_reviewclose(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_reviewclose(uint32_t*, uint32_t*, bool*);

void launch_plan_convert(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_convert(uint32_t*, uint32_t*, bool*);
void launch_plan_tensorboard(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_tensorboard(uint32_t*, uint32_t*, bool*);
void launch_plan_pvcviewer(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_pvcviewer(uint32_t*, uint32_t*, bool*);
void launch_plan_pvcviewerdefault(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_pvcviewerdefault(uint32_t*, uint32_t*, bool*);
void launch_plan_tbupdate(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_tbupdate(uint32_t*, uint32_t*, bool*);
void launch_plan_pvstatus(const PlanIn&, const PlanOut&, uint64_t, cudaStream_t);
void gen_shape_pvstatus(uint32_t*, uint32_t*, bool*);

void launch_plan(int gen, const PlanIn& in, const PlanOut& out, uint64_t in_bytes, cudaStream_t st) {
  switch (gen) {
    case GEN_ID_NBC: launch_plan_nbc(in, out, in_bytes, st); break;
    case GEN_ID_ODH: launch_plan_odh(in, out, in_bytes, st); break;
    case GEN_ID_COPY: launch_plan_copy(in, out, in_bytes, st); break;
    case GEN_ID_STATUS: launch_plan_status(in, out, in_bytes, st); break;
    case GEN_ID_CULL: launch_plan_cull(in, out, in_bytes, st); break;
    case GEN_ID_MUTATE: launch_plan_mutate(in, out, in_bytes, st); break;
    case GEN_ID_GUARD: launch_plan_guard(in, out, in_bytes, st); break;
    case GEN_ID_PODDEFAULT: launch_plan_poddefault(in, out, in_bytes, st); break;
    case GEN_ID_PROFILE: launch_plan_profile(in, out, in_bytes, st); break;
    case GEN_ID_ODHDECIDE: launch_plan_odhdecide(in, out, in_bytes, st); break;
    case GEN_ID_DIFF: launch_plan_diff(in, out, in_bytes, st); break;
    case GEN_ID_REVIEW_OPEN: launch_plan_reviewopen(in, out, in_bytes, st); break;
    case GEN_ID_REVIEW_CLOSE: launch_plan_reviewclose(in, out, in_bytes, st); break;
    case GEN_ID_CONVERT: launch_plan_convert(in, out, in_bytes, st); break;
    case GEN_ID_TENSORBOARD: launch_plan_tensorboard(in, out, in_bytes, st); break;
    case GEN_ID_PVCVIEWER: launch_plan_pvcviewer(in, out, in_bytes, st); break;
    case GEN_ID_PVCVIEWER_DEFAULT: launch_plan_pvcviewerdefault(in, out, in_bytes, st); break;
    case GEN_ID_TB_UPDATE: launch_plan_tbupdate(in, out, in_bytes, st); break;
    case GEN_ID_PVCVIEWER_STATUS: launch_plan_pvstatus(in, out, in_bytes, st); break;
  }
}

static bool shape_of(int gen, uint32_t* nobj, uint32_t* cap, bool* shared1) {
  switch (gen) {
    case GEN_ID_NBC: gen_shape_nbc(nobj, cap, shared1); return true;
    case GEN_ID_ODH: gen_shape_odh(nobj, cap, shared1); return true;
    case GEN_ID_COPY: gen_shape_copy(nobj, cap, shared1); return true;
    case GEN_ID_STATUS: gen_shape_status(nobj, cap, shared1); return true;
    case GEN_ID_CULL: gen_shape_cull(nobj, cap, shared1); return true;
    case GEN_ID_MUTATE: gen_shape_mutate(nobj, cap, shared1); return true;
    case GEN_ID_GUARD: gen_shape_guard(nobj, cap, shared1); return true;
    case GEN_ID_PODDEFAULT: gen_shape_poddefault(nobj, cap, shared1); return true;
    case GEN_ID_PROFILE: gen_shape_profile(nobj, cap, shared1); return true;
    case GEN_ID_ODHDECIDE: gen_shape_odhdecide(nobj, cap, shared1); return true;
    case GEN_ID_DIFF: gen_shape_diff(nobj, cap, shared1); return true;
    case GEN_ID_REVIEW_OPEN: gen_shape_reviewopen(nobj, cap, shared1); return true;
    case GEN_ID_REVIEW_CLOSE: gen_shape_reviewclose(nobj, cap, shared1); return true;
    case GEN_ID_CONVERT: gen_shape_convert(nobj, cap, shared1); return true;
    case GEN_ID_TENSORBOARD: gen_shape_tensorboard(nobj, cap, shared1); return true;
    case GEN_ID_PVCVIEWER: gen_shape_pvcviewer(nobj, cap, shared1); return true;
    case GEN_ID_PVCVIEWER_DEFAULT: gen_shape_pvcviewerdefault(nobj, cap, shared1); return true;
    case GEN_ID_TB_UPDATE: gen_shape_tbupdate(nobj, cap, shared1); return true;
    case GEN_ID_PVCVIEWER_STATUS: gen_shape_pvstatus(nobj, cap, shared1); return true;
  }
  *nobj = 0;
  *cap = 0;
  *shared1 = false;
  return false;
}

void gen_shape(int gen, uint32_t* nobj, uint32_t* cap) {
  bool s;
  shape_of(gen, nobj, cap, &s);
}

bool gen_views(int gen) { return gen == GEN_ID_NBC || gen == GEN_ID_ODH || gen == GEN_ID_MUTATE; }

bool gen_shared1(int gen) {
  uint32_t a, b;
  bool s;
  shape_of(gen, &a, &b, &s);
  return s;
}

}  // namespace nbrm
