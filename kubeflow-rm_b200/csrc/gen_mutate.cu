// gen_mutate.cu — plan kernels of GenMutate, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_TAPE 1       // planners hop through the structural index of a document that has one (nbrm_tape.cuh)
#define NBRM_WARP 1       // small batches: one tuple per warp (plan_kernel_warp)
#define NBRM_WIDE_SKIP 2  // text walk: 8-byte-word container skip (nbrm_json.cuh; measured in profiles/README.md)
#define NBRM_NO_STAGING 1  // the walk reads global memory: staging 128 requests in 100 KB of shared memory left 8 warps per SM (0.72 -> 0.48 ms)
#include "nbrm_mutate.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(mutate, GenMutate)
}  // namespace nbrm
