// gen_status.cu — plan kernel of GenStatus, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_WIDE_SKIP 2  // container skip: 8-byte-word scanner inlined (nbrm_json.cuh; measured in profiles/README.md)
#include "nbrm_status.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(status, GenStatus)
}  // namespace nbrm
