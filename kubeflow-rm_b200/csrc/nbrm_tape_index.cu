// nbrm_tape_index.cu — the index pass of nbrm_tape.cuh as kernels of its own.
//
// index_slot_kernel (nbrm_slot_index): indexes every document of an input slot once and
// keeps the result in HBM beside the slot, for documents that stay resident while many
// batches are planned against them — the found objects, StatefulSets and Pods of the
// cluster state, the informer cache of the reference.  The plan kernels then hop through
// that index instead of walking those documents' text on every pass.
//
// tape_dump_kernel (nbrm_tape_dump): the same pass copied out for inspection; tests
// compare it with the scalar builder entry by entry.
#define NBRM_TAPE 1
#include "nbrm_json.cuh"
#include "nbrm_dev.h"

namespace nbrm {

__global__ void __launch_bounds__(32) tape_dump_kernel(DevDocs docs, uint32_t n, uint32_t cap, TapeEnt* out,
                                                       uint32_t* out_nt, uint32_t* out_flags) {
  extern __shared__ __align__(16) char dump_smem[];
  TapeEnt* tape = reinterpret_cast<TapeEnt*>(dump_smem);
  TapeScratch* sc = reinterpret_cast<TapeScratch*>(dump_smem + (size_t)cap * sizeof(TapeEnt));
  const uint32_t i = blockIdx.x;
  if (i >= n) return;
  const uint64_t o0 = docs.off[i], o1 = docs.off[i + 1];
  uint32_t tf = 0;
  const uint32_t nt = build_tape_warp(docs.data, o0, (uint32_t)(o1 - o0), tape, cap, sc, TAPE_MAX_DEPTH, &tf);
  __syncwarp();
  for (uint32_t k = threadIdx.x; k < nt; k += 32) out[(size_t)i * cap + k] = tape[k];
  if (threadIdx.x == 0) {
    out_nt[i] = nt;
    out_flags[i] = tf;
  }
}

// One warp per document: build in shared memory, then claim room in the slot's tape arena
// and copy.  A document whose index does not fit (the warp's shared memory, or the arena)
// gets none and is walked as text.
static const uint32_t INDEX_WARPS = 4;

__global__ void __launch_bounds__(INDEX_WARPS * 32) index_slot_kernel(DevDocs docs, uint32_t n, uint32_t cap,
                                                                      TapeEnt* arena, uint32_t arena_cap,
                                                                      uint32_t* cursor, uint2* tidx) {
  extern __shared__ __align__(16) char index_smem[];
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t i = blockIdx.x * INDEX_WARPS + warp;
  if (i >= n) return;
  TapeEnt* tape = reinterpret_cast<TapeEnt*>(index_smem) + (size_t)warp * cap;
  TapeScratch* sc = reinterpret_cast<TapeScratch*>(index_smem + (size_t)INDEX_WARPS * cap * sizeof(TapeEnt)) + warp;
  const uint64_t o0 = docs.off[i], o1 = docs.off[i + 1];
  uint32_t tf = DOC_WS;
  uint32_t nt = build_tape_warp(docs.data, o0, (uint32_t)(o1 - o0), tape, cap, sc, TAPE_MAX_DEPTH, &tf);
  __syncwarp();
  uint32_t at = 0;
  if (lane == 0 && nt) at = atomicAdd(cursor, nt);
  at = __shfl_sync(0xffffffffu, at, 0);
  if (nt && (at > arena_cap || nt > arena_cap - at)) nt = 0;  // the arena is full
  for (uint32_t k = lane; k < nt; k += 32) arena[at + k] = tape[k];
  if (lane == 0) tidx[i] = make_uint2(at, nt | (tf << 16));
}

void launch_index_slot(const DevDocs& docs, uint32_t n, void* arena, uint32_t arena_cap, uint32_t* cursor,
                       void* tidx, cudaStream_t st) {
  if (!n) return;
  const uint32_t cap = 2048;
  const size_t smem = (size_t)INDEX_WARPS * (cap * sizeof(TapeEnt) + sizeof(TapeScratch));
  cudaFuncSetAttribute(index_slot_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaMemsetAsync(cursor, 0, sizeof(uint32_t), st);
  index_slot_kernel<<<(n + INDEX_WARPS - 1) / INDEX_WARPS, INDEX_WARPS * 32, smem, st>>>(
      docs, n, cap, reinterpret_cast<TapeEnt*>(arena), arena_cap, cursor, reinterpret_cast<uint2*>(tidx));
}

void launch_tape_dump(const DevDocs& docs, uint32_t n, uint32_t cap, void* out, uint32_t* out_nt,
                      uint32_t* out_flags, cudaStream_t st) {
  if (!n) return;
  const size_t smem = (size_t)cap * sizeof(TapeEnt) + sizeof(TapeScratch);
  cudaFuncSetAttribute(tape_dump_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  tape_dump_kernel<<<n, 32, smem, st>>>(docs, n, cap, reinterpret_cast<TapeEnt*>(out), out_nt, out_flags);
}

}  // namespace nbrm
