// gen_odh.cu — plan kernels of GenOdh, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_TAPE 1       // planners hop through the structural index of a document that has one (nbrm_tape.cuh)
#define NBRM_WARP 1       // small batches: one tuple per warp (plan_kernel_warp)
#define NBRM_WIDE_SKIP 1  // text walk: 8-byte-word container skip (nbrm_json.cuh; measured in profiles/README.md)
#define NBRM_BYTEWISE_LITERALS 1  // literal compares stay byte loops here: the 8-byte form cost this kernel 0.53 -> 0.73 ms
#include "nbrm_odh.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(odh, GenOdh)
}  // namespace nbrm
