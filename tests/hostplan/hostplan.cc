// hostplan.cc — TEST INFRASTRUCTURE.  Compiles the planner headers of libnbrm
// (kubeflow-rm_b200/csrc/*.cuh, which are __host__ __device__) with g++ and
// runs plan + a scalar copy on the CPU, so the walker logic can be checked
// against the oracle in the GPU-less container.  Never loaded by the product
// path; the product is the CUDA build and has no CPU fallback.
#include <stdint.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/nbrm.h"
#include "../../kubeflow-rm_b200/csrc/nbrm_nbc.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_pool.h"

using namespace nbrm;

namespace {
const uint32_t AUX_PAD = 128;

struct Out {
  std::string data;
  std::vector<uint64_t> off;
  std::vector<uint32_t> status;
};
Out g_out;

void emit(const std::vector<Seg>& segs, uint32_t n, const char* const bases[SRC_COUNT],
          std::string& out) {
  for (uint32_t k = 0; k < n; ++k) {
    uint32_t len = segs[k].len_src & SEG_LEN_MASK;
    out.append(bases[segs[k].len_src >> 28] + segs[k].off, len);
  }
}
}  // namespace

extern "C" {

int hostplan_nbc_generate(const char* data, const uint64_t* off, uint64_t n, const nbrm_env* env,
                          uint32_t flags, uint32_t cap, const char** out_data,
                          const uint64_t** out_off, const uint32_t** out_status) {
  std::string pool;
  EnvLits L;
  build_pool_host(env, pool, L);
  uint64_t total_in = n ? off[n] : 0;
  std::vector<char> aux(total_in + n * AUX_PAD + 16);
  g_out.data.clear();
  g_out.off.assign(1, 0);
  g_out.status.assign(n, 0);
  std::vector<Seg> segs(cap);
  const char* bases[SRC_COUNT] = {data, nullptr, nullptr, pool.data(), aux.data()};
  for (uint64_t i = 0; i < n; ++i) {
    Doc d;
    d.p = data + off[i];
    d.n = (uint32_t)(off[i + 1] - off[i]);
    NbView v = view_notebook(d);
    uint32_t st = v.err;
    if (!st && (!is_object(d, v.podspec) || !v.c0.ok())) st = E_NO_CONTAINERS;
    Aux ax;
    ax.abs = (uint32_t)off[i] + (uint32_t)i * AUX_PAD;
    ax.p = aux.data() + ax.abs;
    ax.cap = d.n + AUX_PAD;
    ax.used = 0;
    std::string objs[3];
    for (uint32_t k = 0; k < 3; ++k) {
      Planner P;
      P.base[0] = (uint32_t)off[i];
      P.base[1] = P.base[2] = 0;
      P.begin(segs.data(), cap);
      P.err = E_OK;
      if (!st && (flags & (1u << k))) {
        if (k == 0) plan_statefulset(P, d, v, L, flags);
        else if (k == 1) plan_service(P, d, v, flags);
        else plan_virtualservice(P, d, v, L, flags, ax);
        if (P.err) st = P.err;
        else {
          emit(segs, P.n, bases, objs[k]);
          if (objs[k].size() != P.total) return -3;
        }
      }
    }
    for (uint32_t k = 0; k < 3; ++k) {
      if (!st) g_out.data += objs[k];
      g_out.off.push_back(g_out.data.size());
    }
    g_out.status[i] = st;
  }
  *out_data = g_out.data.data();
  *out_off = g_out.off.data();
  *out_status = g_out.status.data();
  return 0;
}
}
