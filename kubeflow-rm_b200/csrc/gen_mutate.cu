// gen_mutate.cu — plan kernel of GenMutate, in a unit of its own (see nbrm_plan_kernel.cuh).
#include "nbrm_mutate.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(mutate, GenMutate)
}  // namespace nbrm
