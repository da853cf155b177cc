// gen_convert.cu — plan kernel of GenConvert, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_WIDE_SKIP 1  // container skip: one out-of-line 8-byte-word scanner (nbrm_json.cuh)
#include "nbrm_convert.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(convert, GenConvert)
}  // namespace nbrm
