// gen_odhdecide.cu — plan kernel of GenOdhDecide, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_WIDE_SKIP 2  // container skip: 8-byte-word scanner inlined (nbrm_json.cuh)
#include "nbrm_odhcmp.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(odhdecide, GenOdhDecide)
}  // namespace nbrm
