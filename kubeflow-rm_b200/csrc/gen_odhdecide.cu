// gen_odhdecide.cu — plan kernels of GenOdhDecide, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_TAPE 1       // planners hop through the structural index of a document that has one (nbrm_tape.cuh)
#define NBRM_WARP 1       // small batches: one tuple per warp (plan_kernel_warp)
#define NBRM_WIDE_SKIP 1  // text walk: one out-of-line 8-byte-word container skip (inlined, the unit takes 12 minutes to compile)
#define NBRM_WIDE_EQ 16  // spans of at least this many bytes are compared in 8-byte chunks (spans_eq)
#include "nbrm_odhcmp.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(odhdecide, GenOdhDecide)
}  // namespace nbrm
