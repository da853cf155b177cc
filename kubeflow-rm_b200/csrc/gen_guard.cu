// gen_guard.cu — plan kernel of GenRestartGuard, in a unit of its own (see nbrm_plan_kernel.cuh).
#include "nbrm_mutate.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(guard, GenRestartGuard)
}  // namespace nbrm
