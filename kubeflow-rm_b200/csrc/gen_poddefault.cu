// gen_poddefault.cu — plan kernels of GenPodDefault, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_TAPE 1       // planners hop through the structural index of a document that has one (nbrm_tape.cuh)
#define NBRM_WARP 1       // small batches: one tuple per warp (plan_kernel_warp)
#define NBRM_WIDE_SKIP 1  // text walk: one out-of-line 8-byte-word container skip (nbrm_json.cuh)
#include "nbrm_pd.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(poddefault, GenPodDefault)
}  // namespace nbrm
