// gen_cull.cu — plan kernel of GenCull, in a unit of its own (see nbrm_plan_kernel.cuh).
#include "nbrm_cull.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(cull, GenCull)
}  // namespace nbrm
