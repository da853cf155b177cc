// gen_copy.cu — plan kernel of GenCopyFields, in a unit of its own (see nbrm_plan_kernel.cuh).
#include "nbrm_cmp.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(copy, GenCopyFields)
}  // namespace nbrm
