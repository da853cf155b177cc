// gen_odh.cu — plan kernel of GenOdh, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_WIDE_SKIP 1  // container skip: one out-of-line 8-byte-word scanner (nbrm_json.cuh; measured in profiles/README.md)
#include "nbrm_odh.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(odh, GenOdh)
}  // namespace nbrm
