// gen_odh.cu — plan kernel of GenOdh, in a unit of its own (see nbrm_plan_kernel.cuh).
#include "nbrm_odh.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(odh, GenOdh)
}  // namespace nbrm
