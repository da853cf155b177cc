// nbrm_group.cu — order in which a plan kernel's threads take their tuples.
//
// A thread walks one document and the 32 lanes of a warp share every instruction: lanes whose
// documents have the same shape run in lockstep, lanes whose documents differ take turns.  On a
// batch that mixes many shapes (a real cluster: the "hetero" workload of bench.py) the walk
// costs up to four times what it costs on a batch of one shape.  With NBRM_OPT_GROUP_BY_LENGTH
// the tuples are handed to the threads in the order of their first document's length (a stable
// radix sort of the lengths; ties keep their order, so a uniform batch keeps its coalescing):
// documents of one shape have one length and end up in the same warps.  Results stay where
// they were: thread t plans tuple perm[t] and writes to perm[t]'s slots.
#include <cub/device/device_radix_sort.cuh>

#include "nbrm_dev.h"

namespace nbrm {

__global__ void group_keys_kernel(const uint64_t* __restrict__ off, uint32_t n, uint32_t* __restrict__ keys,
                                  uint32_t* __restrict__ vals) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    keys[i] = (uint32_t)(off[i + 1] - off[i]);
    vals[i] = i;
  }
}

size_t group_tmp_bytes(uint32_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, (int)n, 0, 32);
  return bytes + 256;
}

// buf: 4 n words (keys, sorted keys, indices, perm); returns perm.
const uint32_t* launch_group_by_length(const uint64_t* off, uint32_t n, uint32_t* buf, void* tmp, size_t tmp_bytes,
                                       cudaStream_t st) {
  uint32_t *ka = buf, *kb = buf + n, *va = buf + 2 * (size_t)n, *vb = buf + 3 * (size_t)n;
  const uint32_t blocks = (n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184;
  group_keys_kernel<<<blocks, 256, 0, st>>>(off, n, ka, va);
  cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, ka, kb, va, vb, (int)n, 0, 32, st);
  return vb;
}

}  // namespace nbrm
