// gen_profile.cu — plan kernel of GenProfile, in a unit of its own (see nbrm_plan_kernel.cuh).
#include "nbrm_profile.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(profile, GenProfile)
}  // namespace nbrm
