// hostplan.cc — TEST INFRASTRUCTURE.  Compiles the planner headers of libnbrm
// (kubeflow-rm_b200/csrc/*.cuh, which are __host__ __device__) with g++ and
// runs plan + a scalar copy on the CPU, so the walker logic can be checked
// against the oracle in the GPU-less container.  Never loaded by the product
// path; the product is the CUDA build and has no CPU fallback.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "../../include/nbrm.h"
#include "../../kubeflow-rm_b200/csrc/nbrm_nbc.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_odh.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_cmp.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_odhcmp.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_status.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_cull.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_mutate.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_pd.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_profile.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_patch.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_convert.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_tb.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_pvcv.cuh"
#include "../../kubeflow-rm_b200/csrc/nbrm_pool.h"
#include "../../kubeflow-rm_b200/csrc/nbrm_emit.cuh"

using namespace nbrm;

namespace {
const uint32_t AUX_PAD = 128;

struct Out {
  std::string data;
  std::vector<uint64_t> off;
  std::vector<uint32_t> status;
  std::vector<uint32_t> result;
  std::vector<ViewRec> views;  // NOBJ per tuple, as the generator recorded them
};
Out g_out;
#ifdef NBRM_TAPE
uint32_t g_tape_cap = 65000;  // entries per document; hostplan_set_tape_cap shrinks it to test the fallback
#endif

template <class G>
int run_host(Out& g_out, const char* const data[3], const uint64_t* const off[3], uint64_t n, const nbrm_env* env,
             uint32_t flags, const void* extra, uint64_t now, uint32_t arg0, uint32_t arg1,
             uint32_t shared_n, uint64_t first = 0) {
  std::string pool;
  EnvLits L;
  build_pool_host(env, pool, L);
  uint64_t total_in = 0;
  for (int k = 0; k < 3; ++k)
    if (data[k] && n && !(G::SHARED1 && k == 1)) total_in += off[k][n];
  std::vector<char> aux(total_in + n * AUX_PAD + 16);
  g_out.data.clear();
  g_out.off.assign(1, 0);
  g_out.status.assign(n, 0);
  g_out.result.assign(n, 0);
  const uint32_t cap = G::CAP;
  std::vector<Seg> segs((size_t)G::NOBJ * cap);
  std::vector<uint32_t> nseg(G::NOBJ), lens(G::NOBJ);
  std::vector<ViewEnt> views((size_t)G::NOBJ * VIEW_FIELDS);
  g_out.views.clear();
  const char* bases[SRC_COUNT] = {data[0], data[1], data[2], pool.data(), aux.data()};
#ifdef NBRM_TAPE
  // the shared arena (PodDefault list) is indexed once, as nbrm_slot_index does on the device
  std::vector<TapeEnt> sh_tape;
  std::vector<uint32_t> sh_tidx;
  if (G::SHARED1 && data[1]) {
    for (uint32_t j = 0; j < shared_n; ++j) {
      const uint32_t len = (uint32_t)(off[1][j + 1] - off[1][j]);
      const size_t at = sh_tape.size();
      sh_tape.resize(at + len + 1);
      uint32_t tf = 0;
      const uint32_t nt = build_tape_scalar(data[1] + off[1][j], len, sh_tape.data() + at, g_tape_cap, &tf);
      sh_tape.resize(at + nt);
      sh_tidx.push_back((uint32_t)at);
      sh_tidx.push_back(nt | (tf << 16));
    }
  }
#endif
  for (uint64_t i = 0; i < n; ++i) {
    DocCtx c;
    uint32_t aux_abs = (uint32_t)i * AUX_PAD, aux_cap = AUX_PAD;
#ifdef NBRM_TAPE
    std::vector<TapeEnt> tapes[3];
#endif
    for (int k = 0; k < 3; ++k) {
      if (data[k] && !(G::SHARED1 && k == 1)) {
        c.d[k].p = data[k] + off[k][i];
        c.d[k].n = (uint32_t)(off[k][i + 1] - off[k][i]);
        c.base[k] = (uint32_t)off[k][i];
        aux_abs += (uint32_t)off[k][i];
        aux_cap += c.d[k].n;
#ifdef NBRM_TAPE
        // tape mode: the scalar builder stands in for the warp's index pass
        tapes[k].resize(c.d[k].n + 1);
        c.d[k].nt = build_tape_scalar(c.d[k].p, c.d[k].n, tapes[k].data(), g_tape_cap, &c.d[k].tflags);
        c.d[k].tape = tapes[k].data();
        c.d[k].dlim = TAPE_MAX_DEPTH;
#endif
      } else {
        c.d[k].p = nullptr;
        c.d[k].n = 0;
        c.base[k] = 0;
      }
    }
    c.aux.abs = aux_abs;
    c.aux.p = aux.data() + aux_abs;
    c.aux.cap = aux_cap;
    c.aux.used = 0;
    c.env = &L;
    c.flags = flags;
    c.index = (uint32_t)(first + i);
    c.extra = extra;
    c.now = now;
    c.arg0 = arg0;
    c.arg1 = arg1;
    c.result = 0;
    c.shared = G::SHARED1 ? data[1] : nullptr;
    c.shared_off = G::SHARED1 ? off[1] : nullptr;
    c.shared_n = G::SHARED1 ? shared_n : 0;
#ifdef NBRM_TAPE
    if (G::SHARED1 && data[1]) {
      c.shared_tape = sh_tape.data();
      c.shared_tidx = sh_tidx.data();
    }
#endif
    ObjSink s;
    s.segs = segs.data();
    s.nseg = nseg.data();
    s.lens = lens.data();
    s.cap = cap;
    memset(views.data(), 0, views.size() * sizeof(ViewEnt));  // as the host zeroes the table before a plan kernel
    s.views.p = views.data();
    s.views.n = G::NOBJ;
    s.view_j = 0;
    uint32_t st = G::run(c, s);
    for (uint32_t k = 0; k < G::NOBJ; ++k) {  // back to one record per object for the tests
      ViewRec r;
      for (uint32_t f = 0; f < VIEW_FIELDS; ++f) {
        r.off[f] = views[(size_t)f * G::NOBJ + k].off;
        r.len[f] = views[(size_t)f * G::NOBJ + k].len;
      }
      g_out.views.push_back(r);
    }
    g_out.status[i] = st;
    g_out.result[i] = c.result;
    for (uint32_t k = 0; k < G::NOBJ; ++k) {
      size_t before = g_out.data.size();
      for (uint32_t j = 0; j < nseg[k]; ++j) {
        const Seg& sg = segs[(size_t)k * cap + j];
        g_out.data.append(bases[sg.len_src >> 28] + sg.off, sg.len_src & SEG_LEN_MASK);
      }
      if (g_out.data.size() - before != lens[k]) return -3;  // planned length must equal emitted bytes
      g_out.off.push_back(g_out.data.size());
    }
  }
  return 0;
}
}  // namespace

#define HP_API __attribute__((visibility("default")))

extern "C" {

// The emit kernel's per-block routine (nbrm_emit.cuh), run block by block the way emit_warp's
// rounds hand blocks to lanes.  srcs[k] / src_len[k]: the source buffers; seg_*: the segment
// list of one object; head: where in a 16-byte block of the output the object starts; win:
// segments per window.  out must hold head + total + 16 bytes and comes back with every
// byte outside [head, head + total) untouched.
HP_API int hostplan_emit_blocks(const char* const* srcs, const uint64_t* src_len, uint32_t n_src,
                                const uint32_t* seg_off, const uint32_t* seg_len, const uint32_t* seg_src,
                                uint32_t ns, uint32_t head, uint32_t win, char* out) {
  if (n_src > SRC_COUNT || head > 15 || !win) return -1;
  // sources in 16-byte aligned, padded copies: the routine reads whole aligned blocks
  std::vector<std::vector<char>> copies(n_src);
  const char* bases[SRC_COUNT] = {};
  for (uint32_t k = 0; k < n_src; ++k) {
    copies[k].assign(src_len[k] + 48, 0x55);
    char* a = copies[k].data();
    a += (16 - ((uintptr_t)a & 15)) & 15;
    memcpy(a, srcs[k], src_len[k]);
    bases[k] = a;
  }
  std::vector<char> obuf(head + 64, 0);
  uint64_t total = 0;
  for (uint32_t j = 0; j < ns; ++j) total += seg_len[j];
  obuf.assign(head + total + 64, 0);
  char* o = obuf.data();
  o += (16 - ((uintptr_t)o & 15)) & 15;
  memcpy(o, out, head + total + 16);
  char* dst = o + head;
  for (uint32_t w0 = 0; w0 < ns; w0 += win) {
    const uint32_t m = ns - w0 < win ? ns - w0 : win;
    std::vector<uint32_t> P(m + 1, 0);
    std::vector<Seg> S(m);
    for (uint32_t j = 0; j < m; ++j) {
      S[j].off = seg_off[w0 + j];
      S[j].len_src = (seg_src[w0 + j] << 28) | seg_len[w0 + j];
      P[j + 1] = P[j] + seg_len[w0 + j];
    }
    const uint32_t L = P[m];
    const uint32_t hd = (uint32_t)((uintptr_t)dst & 15);
    char* blk0 = dst - hd;
    const uint32_t nblocks = (hd + L + 15) >> 4;
    uint32_t first = 0;
    uint32_t T[17 * 4];
    for (uint32_t x = 0; x < 17; ++x)
      for (uint32_t w = 0; w < 4; ++w) T[4 * x + w] = emitw::below_word(x, w);
    for (uint32_t c = 0; c < nblocks; c += 32) {  // the rounds of emit_warp, lanes one after the other
      const uint32_t nbr = std::min<uint32_t>(32, nblocks - c);
      const uint32_t qs = c ? 16 * c - hd : 0;
      const uint32_t qe = std::min<uint32_t>(L, 16 * (c + nbr) - hd);
      uint32_t cnt = 0;
      while (first + 1 + cnt < m && P[first + 1 + cnt] < qe) ++cnt;
      const bool ragged = (c == 0 && hd) || 16 * (c + nbr) - hd > L;
      char* rdst = blk0 + 16 * (size_t)c;
      for (uint32_t lane = 0; lane < nbr; ++lane) {
        if (cnt == 0 && !ragged) {
          emitw::copy_pure_block(bases[S[first].len_src >> 28] + S[first].off + (qs - P[first]), rdst, lane);
          continue;
        }
        const uint32_t b = c + lane;
        const uint32_t q = b ? 16 * b - hd : 0;
        const uint32_t g = cnt ? emitw::find_seg(P.data(), first, first + cnt, q) : first;
        uint32_t X[12];
        memset(X, 0xcc, sizeof X);
        emitw::emit_block(P.data(), S.data(), g, bases, q, std::min<uint32_t>(L, 16 * (b + 1) - hd), b ? 0 : hd,
                          rdst + 16 * (size_t)lane, X, T);
      }
      first += cnt;
      if (first + 1 < m && P[first + 1] == qe) ++first;
    }
    dst += L;
  }
  memcpy(out, o, head + total + 16);
  return 0;
}

HP_API int hostplan_tape_mode(void) {
#ifdef NBRM_TAPE
  return 1;
#else
  return 0;
#endif
}

HP_API void hostplan_set_tape_cap(uint32_t cap) {
#ifdef NBRM_TAPE
  g_tape_cap = cap;
#else
  (void)cap;
#endif
}

// The views the last hostplan_run's generator recorded (VIEW_FIELDS offsets then VIEW_FIELDS
// lengths per object), and what view_object finds by walking a document, in the same layout.
HP_API const uint32_t* hostplan_views(uint32_t* fields) {
  *fields = VIEW_FIELDS;
  return reinterpret_cast<const uint32_t*>(g_out.views.data());
}
// A Notebook's view built from a recorded ViewRec (as the generators that read a mutated
// Notebook get it) against the view the full walk finds: 0 when every span and flag agrees.
HP_API int hostplan_check_notebook_view(const char* p, uint32_t n, const uint32_t* rec) {
  Doc d;
  d.p = p;
  d.n = n;
  ViewEnt tab[VIEW_FIELDS];
  for (uint32_t f = 0; f < VIEW_FIELDS; ++f) {
    tab[f].off = rec[f];
    tab[f].len = rec[VIEW_FIELDS + f];
  }
  ViewIn vin;
  vin.t.p = tab;
  vin.t.n = 1;
  vin.j = 0;
  const NbView a = view_notebook(d, vin), b = view_notebook(d);
  auto same = [](Span x, Span y) { return x.len == y.len && (x.len == 0 || x.off == y.off); };
  if (a.err != b.err) return 1;
  if (a.err) return 0;
  if (!same(a.root, b.root)) return 2;
  if (!same(a.meta, b.meta)) return 3;
  if (!same(a.name, b.name) || !same(a.ns, b.ns) || !same(a.uid, b.uid)) return 4;
  if (!same(a.labels, b.labels)) return 5;
  if (!same(a.annotations, b.annotations)) return 6;
  if (!same(a.podspec, b.podspec)) return 7;
  if (!same(a.containers, b.containers)) return 8;
  if (!same(a.c0, b.c0)) return 9;
  if (!same(a.a_stopped, b.a_stopped) || !same(a.a_rewrite, b.a_rewrite) || !same(a.a_headers, b.a_headers) ||
      !same(a.a_oauth, b.a_oauth) || !same(a.a_mesh, b.a_mesh) || !same(a.a_logout, b.a_logout) ||
      !same(a.a_restart, b.a_restart))
    return 10;
  if (a.l_statefulset != b.l_statefulset || a.l_notebook_name != b.l_notebook_name || a.l_workbench != b.l_workbench)
    return 11;
  // 100: the recorded extents were complete enough for the short walk to be the one that ran
  return (rec[VIEW_FIELDS + VF_ROOT] && rec[VIEW_FIELDS + VF_META] && rec[VIEW_FIELDS + VF_TPOD] &&
          rec[VIEW_FIELDS + VF_PORTS] && rec[VIEW_FIELDS + VF_SUBJECTS])
             ? 100
             : 0;
}

HP_API void hostplan_view_object(const char* p, uint32_t n, uint32_t* out) {
  Doc d;
  d.p = p;
  d.n = n;
  const ObjView v = view_object(d);
  const Span f[VIEW_FIELDS] = {v.root, v.kind, v.meta, v.labels, v.annotations, v.spec, v.replicas,
                               v.tlabels, v.tpod, v.selector, v.ports, find(d, v.root, "subjects")};
  for (uint32_t k = 0; k < VIEW_FIELDS; ++k) {
    out[k] = f[k].off;
    out[VIEW_FIELDS + k] = f[k].len;
  }
}

// The full reconcile pass of kubeflow-rm_b200/reconcile.py (webhook -> ODH objects -> ODH
// write decision -> status -> notebook-controller objects -> create-or-update decision) on
// the CPU, one shard of Notebooks per thread: the native CPU baseline bench.py reports beside
// the Python oracle.  Inputs are packed documents: n requests, 3n found objects, n
// StatefulSets, n Pods, 7n found ODH objects, 16n seed bytes.  Returns the bytes of every
// result (mutated Notebooks, ODH write set, statuses, write set) in out_bytes[4], and the
// number of tuples any stage rejected.
namespace {
struct Packed {
  std::string data;
  std::vector<uint64_t> off;
};
void take(Packed& p, const Out& o) {
  p.data = o.data;
  p.off = o.off;
}
struct PassShard {
  uint64_t bytes[4] = {0, 0, 0, 0};
  uint64_t rejected = 0;
};
void pass_shard(const char* const d[5], const uint64_t* const o[5], const unsigned char* seeds, uint64_t lo,
                uint64_t hi, const nbrm_env* env, uint64_t now, uint32_t nbc_flags, uint32_t odh_flags,
                PassShard* res) {
  const uint64_t n = hi - lo;
  if (!n) return;
  Out out;
  // rebase a slice [a, b) of packed documents to offsets that start at 0
  auto slice = [](const char* data, const uint64_t* off, uint64_t a, uint64_t b, Packed& p) {
    p.data.assign(data + off[a], off[b] - off[a]);
    p.off.resize(b - a + 1);
    for (uint64_t k = a; k <= b; ++k) p.off[k - a] = off[k] - off[a];
  };
  Packed req, found, sts, pod, ofound, nbs, odh, desired;
  slice(d[0], o[0], lo, hi, req);
  slice(d[1], o[1], 3 * lo, 3 * hi, found);
  slice(d[2], o[2], lo, hi, sts);
  slice(d[3], o[3], lo, hi, pod);
  slice(d[4], o[4], 7 * lo, 7 * hi, ofound);
  auto count = [&](const Out& x) {
    for (uint32_t v : x.status) res->rejected += v != 0;
  };
  {  // NotebookWebhook.Handle (CREATE)
    const char* dd[3] = {req.data.data(), nullptr, nullptr};
    const uint64_t* oo[3] = {req.off.data(), nullptr, nullptr};
    run_host<GenMutate>(out, dd, oo, n, env, 1u, nullptr, 0, 0, 0, 0);
    count(out);
    res->bytes[0] += out.data.size();
    take(nbs, out);
  }
  {  // ODH desired objects, then what the ODH reconcilers write
    const char* dd[3] = {nbs.data.data(), nullptr, nullptr};
    const uint64_t* oo[3] = {nbs.off.data(), nullptr, nullptr};
    run_host<GenOdh>(out, dd, oo, n, env, odh_flags, seeds, 0, 0, 0, 0);  // seeds already point at this shard
    count(out);
    take(odh, out);
    const char* d2[3] = {odh.data.data(), ofound.data.data(), nullptr};
    const uint64_t* o2[3] = {odh.off.data(), ofound.off.data(), nullptr};
    run_host<GenOdhDecide>(out, d2, o2, 7 * n, env, 0, nullptr, 0, 0, 0, 0);
    count(out);
    res->bytes[1] += out.data.size();
  }
  {  // createNotebookStatus
    const char* dd[3] = {nbs.data.data(), sts.data.data(), pod.data.data()};
    const uint64_t* oo[3] = {nbs.off.data(), sts.off.data(), pod.off.data()};
    run_host<GenStatus>(out, dd, oo, n, env, 0, nullptr, now, 0, 0, 0);
    count(out);
    res->bytes[2] += out.data.size();
  }
  {  // generateStatefulSet / Service / VirtualService, then Copy*Fields against the found objects
    const char* dd[3] = {nbs.data.data(), nullptr, nullptr};
    const uint64_t* oo[3] = {nbs.off.data(), nullptr, nullptr};
    run_host<GenNbc>(out, dd, oo, n, env, nbc_flags, nullptr, 0, 0, 0, 0);
    count(out);
    take(desired, out);
    const char* d2[3] = {desired.data.data(), found.data.data(), nullptr};
    const uint64_t* o2[3] = {desired.off.data(), found.off.data(), nullptr};
    run_host<GenCopyFields>(out, d2, o2, 3 * n, env, 0, nullptr, 0, 0, 0, 0);
    count(out);
    res->bytes[3] += out.data.size();
  }
}
}  // namespace

HP_API int hostplan_full_pass(const char* req, const uint64_t* req_off, const char* found, const uint64_t* found_off,
                              const char* sts, const uint64_t* sts_off, const char* pod, const uint64_t* pod_off,
                              const char* ofound, const uint64_t* ofound_off, const unsigned char* seeds,
                              uint64_t n, const nbrm_env* env, uint64_t now, uint32_t nbc_flags,
                              uint32_t odh_flags, uint32_t threads, uint64_t* out_bytes, uint64_t* rejected) {
  if (threads < 1) threads = 1;
  const char* d[5] = {req, found, sts, pod, ofound};
  const uint64_t* o[5] = {req_off, found_off, sts_off, pod_off, ofound_off};
  std::vector<PassShard> res(threads);
  std::vector<std::thread> pool;
  for (uint32_t t = 0; t < threads; ++t) {
    const uint64_t lo = n * t / threads, hi = n * (t + 1) / threads;
    pool.emplace_back(pass_shard, d, o, seeds + 16 * lo, lo, hi, env, now, nbc_flags, odh_flags, &res[t]);
  }
  for (auto& th : pool) th.join();
  for (int k = 0; k < 4; ++k) out_bytes[k] = 0;
  *rejected = 0;
  for (auto& r : res) {
    for (int k = 0; k < 4; ++k) out_bytes[k] += r.bytes[k];
    *rejected += r.rejected;
  }
  return 0;
}

// The scalar index builder on one document (tape build only); returns the entry count.
HP_API uint32_t hostplan_build_tape(const char* p, uint32_t n, uint64_t* out, uint32_t cap, uint32_t* flags) {
#ifdef NBRM_TAPE
  static_assert(sizeof(TapeEnt) == 8, "entries are 64-bit words");
  return build_tape_scalar(p, n, reinterpret_cast<TapeEnt*>(out), cap, flags);
#else
  (void)p; (void)n; (void)out; (void)cap; (void)flags;
  return 0;
#endif
}

// gen: 0 notebook-controller (3 objects), 1 ODH reconcile (7), 2 Copy*Fields (1), 3 status (1), 4 culling (1)
HP_API int hostplan_run(int gen, const char* d0, const uint64_t* o0, const char* d1, const uint64_t* o1,
                 const char* d2, const uint64_t* o2, uint64_t n, const nbrm_env* env, uint32_t flags,
                 const void* extra, uint64_t now, uint32_t arg0, uint32_t arg1, uint32_t shared_n,
                 uint32_t* nobj,
                 const char** out_data, const uint64_t** out_off, const uint32_t** out_status,
                 const uint32_t** out_result) {
  const char* data[3] = {d0, d1, d2};
  const uint64_t* off[3] = {o0, o1, o2};
  int rc = -9;
  switch (gen) {
    case 0: rc = run_host<GenNbc>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenNbc::NOBJ; break;
    case 1: rc = run_host<GenOdh>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenOdh::NOBJ; break;
    case 2: rc = run_host<GenCopyFields>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenCopyFields::NOBJ; break;
    case 3: rc = run_host<GenStatus>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenStatus::NOBJ; break;
    case 4: rc = run_host<GenCull>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenCull::NOBJ; break;
    case 5: rc = run_host<GenMutate>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenMutate::NOBJ; break;
    case 6: rc = run_host<GenRestartGuard>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenRestartGuard::NOBJ; break;
    case 7: rc = run_host<GenPodDefault>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenPodDefault::NOBJ; break;
    case 9: rc = run_host<GenOdhDecide>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenOdhDecide::NOBJ; break;
    case 10: rc = run_host<GenJsonDiff>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenJsonDiff::NOBJ; break;
    case 11: rc = run_host<GenReviewOpen>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenReviewOpen::NOBJ; break;
    case 12: rc = run_host<GenReviewClose>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenReviewClose::NOBJ; break;
    case 13: rc = run_host<GenConvert>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenConvert::NOBJ; break;
    case 17: rc = run_host<GenTbUpdate>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenTbUpdate::NOBJ; break;
    case 18: rc = run_host<GenPvStatus>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenPvStatus::NOBJ; break;
    case 16: rc = run_host<GenPvcViewerDefault>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenPvcViewerDefault::NOBJ; break;
    case 15: rc = run_host<GenPvcViewer>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenPvcViewer::NOBJ; break;
    case 14: rc = run_host<GenTensorboard>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenTensorboard::NOBJ; break;
    case 8: rc = run_host<GenProfile>(g_out, data, off, n, env, flags, extra, now, arg0, arg1, shared_n); *nobj = GenProfile::NOBJ; break;
  }
  *out_data = g_out.data.data();
  *out_off = g_out.off.data();
  *out_status = g_out.status.data();
  *out_result = g_out.result.data();
  return rc;
}
}
