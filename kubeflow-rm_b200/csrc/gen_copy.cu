// gen_copy.cu — plan kernel of GenCopyFields, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_WIDE_SKIP 2  // container skip: 8-byte-word scanner inlined (nbrm_json.cuh; measured in profiles/README.md)
#define NBRM_WIDE_EQ 1
#include "nbrm_cmp.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(copy, GenCopyFields)
}  // namespace nbrm
