// gen_status.cu — plan kernel of GenStatus, in a unit of its own (see nbrm_plan_kernel.cuh).
#include "nbrm_status.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(status, GenStatus)
}  // namespace nbrm
