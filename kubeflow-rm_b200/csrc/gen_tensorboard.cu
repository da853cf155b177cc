// gen_tensorboard.cu — plan kernel of GenTensorboard, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_WIDE_SKIP 1   // strings, keys and containers in 8-byte chunks (nbrm_json.cuh)
#define NBRM_NO_STAGING 1  // walks global memory
#include "nbrm_tb.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(tensorboard, GenTensorboard)
NBRM_DEFINE_GENERATOR(tbupdate, GenTbUpdate)
}  // namespace nbrm
