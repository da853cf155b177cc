// gen_cull.cu — plan kernel of GenCull, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_WIDE_SKIP 1   // strings, keys and containers in 8-byte chunks (nbrm_json.cuh)
#define NBRM_NO_STAGING 1  // walks global memory: 7.5 -> 4.9 ms per 100 000 Notebooks
#include "nbrm_cull.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(cull, GenCull)
}  // namespace nbrm
