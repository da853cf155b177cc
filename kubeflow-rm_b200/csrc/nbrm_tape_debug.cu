// nbrm_tape_debug.cu — nbrm_tape_dump: the structural index of every document of a slot,
// as the warp builder of nbrm_tape.cuh writes it, copied out for inspection.  The planners
// never call this; tests compare it with the scalar builder entry by entry.
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

void launch_tape_dump(const DevDocs& docs, uint32_t n, uint32_t cap, void* out, uint32_t* out_nt,
                      uint32_t* out_flags, cudaStream_t st) {
  if (!n) return;
  const size_t smem = (size_t)cap * sizeof(TapeEnt) + sizeof(TapeScratch);
  cudaFuncSetAttribute(tape_dump_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  tape_dump_kernel<<<n, 32, smem, st>>>(docs, n, cap, reinterpret_cast<TapeEnt*>(out), out_nt, out_flags);
}

}  // namespace nbrm
