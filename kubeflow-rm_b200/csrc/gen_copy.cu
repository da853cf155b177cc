// gen_copy.cu — plan kernel of GenCopyFields, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_TAPE 1       // planners hop through the structural index (nbrm_tape.cuh) ...
#define NBRM_WARP 1       // ... one tuple per warp (plan_kernel_warp)
#define NBRM_WIDE_SKIP 1  // text walk of a document without a tape: one out-of-line 8-byte-word scanner
#include "nbrm_cmp.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(copy, GenCopyFields)
}  // namespace nbrm
