// nbrm_dev.h — internal declarations shared by the .cu files of libnbrm.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "nbrm_plan.cuh"

namespace nbrm {

static const int EMIT_THREADS = 256;
static const int PLAN_THREADS = 128;
static const uint32_t AUX_PAD = 128;  // aux slice of input doc i: in_len(i) + AUX_PAD bytes

struct SrcBases {
  const char* p[SRC_COUNT];
};

// One batch of JSON documents resident in device memory.
struct DevDocs {
  const char* data;     // null when the slot is not used by the generator
  const uint64_t* off;  // n + 1 offsets into data
  // the slot's structural index when it has one (nbrm_slot_index; nbrm_tape_index.cu):
  // tidx[i] = (first entry of document i in tape, entry count | flags << 16)
  const void* tape;
  const uint2* tidx;
  // the producing generator's views of the documents (field-major ViewTable), or null
  ViewEnt* views;
};

// What a plan kernel reads: tuple i = document i of every slot in use.
struct PlanIn {
  DevDocs slot[3];
  uint32_t n;
  uint32_t flags;
  EnvLits env;
  const void* extra;  // generator-specific side input (device pointer) or null
  uint64_t now;       // caller's clock, Unix seconds
  uint32_t arg0, arg1;
  uint32_t shared_n;  // documents in slot 1 when the generator reads it as a shared arena
  uint32_t period;    // kinds of document repeat with this period in slot 0 (1 = all alike)
  const uint32_t* perm;  // thread t plans tuple perm[t] (nbrm_group.cu), or null: tuple t
};

// Where a plan kernel writes.
struct PlanOut {
  Seg* segs;          // n * NOBJ * cap
  uint32_t* nseg;     // n * NOBJ
  uint32_t* lens;     // n * NOBJ
  uint32_t* status;   // one per input tuple
  uint32_t* result;   // one per input tuple, generator-specific
  char* aux;          // aux arena
  uint32_t cap;
  ViewEnt* views;     // VIEW_FIELDS * n * NOBJ view entries of the planned objects (field-major), or null
};

// nbrm_kernels.cu
void launch_scan(const uint32_t* lens, uint32_t n, uint64_t* tile_tmp, uint64_t* off,
                 cudaStream_t st);
size_t scan_tmp_bytes(uint32_t n);
void launch_store_u64(const uint64_t* src, uint64_t* host_dst, cudaStream_t st);
void launch_copy_u32(const uint32_t* src, uint32_t* dst, uint32_t n, cudaStream_t st);
void launch_merge_status(const uint32_t* gen, uint32_t* pair, uint32_t n_pairs, uint32_t per, cudaStream_t st);
void launch_emit(const Seg* segs, const uint32_t* nseg, const uint32_t* lens, const uint64_t* off,
                 uint32_t n_out, uint32_t cap, uint32_t nobj, const SrcBases& bases, char* out,
                 int sm_count, cudaStream_t st);

// nbrm_group.cu
size_t group_tmp_bytes(uint32_t n);
const uint32_t* launch_group_by_length(const uint64_t* off, uint32_t n, uint32_t* buf, void* tmp, size_t tmp_bytes,
                                       cudaStream_t st);

// nbrm_tape_index.cu
void launch_index_slot(const DevDocs& docs, uint32_t n, void* arena, uint32_t arena_cap, uint32_t* cursor,
                       void* tidx, cudaStream_t st);
void launch_tape_dump(const DevDocs& docs, uint32_t n, uint32_t cap, void* out, uint32_t* out_nt,
                      uint32_t* out_flags, cudaStream_t st);

// nbrm_plan_kernels.cu
enum GenId { GEN_ID_NBC = 0, GEN_ID_ODH = 1, GEN_ID_COPY = 2, GEN_ID_STATUS = 3, GEN_ID_CULL = 4,
             GEN_ID_MUTATE = 5, GEN_ID_GUARD = 6, GEN_ID_PODDEFAULT = 7, GEN_ID_PROFILE = 8, GEN_ID_ODHDECIDE = 9, GEN_ID_DIFF = 10,
             GEN_ID_REVIEW_OPEN = 11, GEN_ID_REVIEW_CLOSE = 12, GEN_ID_CONVERT = 13, GEN_ID_TENSORBOARD = 14, GEN_ID_PVCVIEWER = 15, GEN_ID_PVCVIEWER_DEFAULT = 16, GEN_ID_TB_UPDATE = 17, GEN_ID_PVCVIEWER_STATUS = 18 };
void launch_plan(int gen, const PlanIn& in, const PlanOut& out, uint64_t in_bytes, cudaStream_t st);
void gen_shape(int gen, uint32_t* nobj, uint32_t* cap);
bool gen_shared1(int gen);
bool gen_views(int gen);  // the generator leaves views of its objects (nbrm_plan.cuh)

}  // namespace nbrm
