// gen_pvcviewer.cu — plan kernel of GenPvcViewer, in a unit of its own (see nbrm_plan_kernel.cuh).
#define NBRM_WIDE_SKIP 1   // strings, keys and containers in 8-byte chunks (nbrm_json.cuh)
#define NBRM_NO_STAGING 1  // walks global memory
#include "nbrm_pvcv.cuh"
#include "nbrm_plan_kernel.cuh"

namespace nbrm {
NBRM_DEFINE_GENERATOR(pvcviewer, GenPvcViewer)
NBRM_DEFINE_GENERATOR(pvcviewerdefault, GenPvcViewerDefault)
NBRM_DEFINE_GENERATOR(pvstatus, GenPvStatus)
}  // namespace nbrm
