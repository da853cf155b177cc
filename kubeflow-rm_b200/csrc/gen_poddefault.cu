// gen_poddefault.cu — plan kernel of GenPodDefault, in a unit of its own (see nbrm_plan_kernel.cuh).
#include "nbrm_pd.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(poddefault, GenPodDefault)
}  // namespace nbrm
