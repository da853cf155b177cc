// gen_nbc.cu — plan kernel of GenNbc, in a unit of its own (see nbrm_plan_kernel.cuh).
#include "nbrm_nbc.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(nbc, GenNbc)
}  // namespace nbrm
