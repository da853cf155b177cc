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
  const char* data;
  const uint64_t* off;  // n + 1 offsets into data
  uint32_t n;
};

// Where a plan kernel writes.
struct PlanOut {
  Seg* segs;          // n_out * cap
  uint32_t* nseg;     // n_out
  uint32_t* lens;     // n_out
  uint32_t* status;   // one per input document
  char* aux;          // aux arena
  uint32_t cap;
};

// nbrm_kernels.cu
void launch_scan(const uint32_t* lens, uint32_t n, uint64_t* tile_tmp, uint64_t* off,
                 cudaStream_t st);
size_t scan_tmp_bytes(uint32_t n);
void launch_emit(const Seg* segs, const uint32_t* nseg, const uint64_t* off, uint32_t n_out,
                 uint32_t cap, const SrcBases& bases, char* out, int sm_count, cudaStream_t st);

// nbrm_nbc.cu — 3 output objects per Notebook: StatefulSet, Service, VirtualService
static const uint32_t NBC_OUT_PER_DOC = 3;
static const uint32_t NBC_SEG_CAP = 96;
void launch_nbc_generate_plan(const DevDocs& nb, const EnvLits& env, uint32_t flags,
                              const PlanOut& out, cudaStream_t st);

}  // namespace nbrm
