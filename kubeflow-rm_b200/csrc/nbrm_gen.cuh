// nbrm_gen.cuh — the frame every generator runs in: one input document (or one
// aligned tuple of documents from up to three input arenas) in, a fixed number
// of planned output objects out.  The same frame is instantiated by the plan
// kernel (nbrm_plan_kernels.cu) and by the CPU test build of the planners.
#pragma once
#include "nbrm_plan.cuh"

namespace nbrm {

// Per-document slice of the aux arena: bytes a planner computes.
struct Aux {
  char* p;        // start of this document's slice
  uint32_t abs;   // absolute offset of the slice in the aux arena
  uint32_t cap;
  uint32_t used;
};

// Where the objects planned for ONE input tuple go (already offset to its first object).
struct ObjSink {
  Seg* segs;       // n_obj * cap
  uint32_t* nseg;  // n_obj
  uint32_t* lens;  // n_obj
  uint32_t cap;
  ViewTable views = {nullptr, 0};  // where the generator's views of the planned objects go, when someone asked ...
  uint32_t view_j = 0;             // ... and the index of this tuple's first object in that table
};

// Everything a generator sees for one input tuple.
struct DocCtx {
  Doc d[3];          // document i of input slot 0..2 (n == 0 when the slot is unused or the doc is empty)
  uint32_t base[3];  // its absolute offset in that slot's arena
  Aux aux;
  const EnvLits* env;
  uint32_t flags;
  uint32_t index;       // document index within the batch
  const void* extra;    // generator-specific side input (device pointer), may be null
  uint64_t now;         // caller-supplied clock, Unix seconds (generators that stamp times)
  uint32_t arg0, arg1;  // generator-specific integers (e.g. culling periods)
  // generators with SHARED1: input slot 1 as one arena every tuple may read
  const char* shared;
  const uint64_t* shared_off;
  uint32_t shared_n;
  // ... and its structural index when the arena was indexed (nbrm_slot_index): entries, and per
  // document (first entry, count | flags << 16)
  const void* shared_tape = nullptr;
  const void* shared_tidx = nullptr;
  uint32_t result;      // generator-specific per-tuple result word (0 when unused)
  ViewIn view[3] = {{{nullptr, 0}, 0}, {{nullptr, 0}, 0}, {{nullptr, 0}, 0}};  // the producing generator's view of d[k], when it left one
};

// Plans n_obj objects with plan_k(P, k).  One failing object voids the whole
// tuple: all its objects get length 0 and the error is returned.
template <class PlanK>
NB_HD uint32_t plan_objects(const DocCtx& c, ObjSink s, uint32_t n_obj, uint32_t st, PlanK& plan_k) {
  Planner P;
  P.base[0] = c.base[0];
  P.base[1] = c.base[1];
  P.base[2] = c.base[2];
  for (uint32_t k = 0; k < n_obj; ++k) {
    P.begin(s.segs + (size_t)k * s.cap, s.cap);
    P.err = E_OK;
    if (!st) {
      plan_k(P, k);
      P.finish();
      if (P.err) st = P.err;
    }
    s.nseg[k] = st ? 0 : P.n;
    s.lens[k] = st ? 0 : P.total;
  }
  if (st) {
    for (uint32_t k = 0; k < n_obj; ++k) {
      s.nseg[k] = 0;
      s.lens[k] = 0;
    }
  }
  return st;
}

}  // namespace nbrm
