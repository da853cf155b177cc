// nbrm_capi.cu — the extern "C" surface declared in include/nbrm.h.
#include <stdio.h>
#include <string.h>

#include <string>
#include <utility>
#include <vector>

#include "../../include/nbrm.h"
#include "nbrm_dev.h"
#include "nbrm_pool.h"

using namespace nbrm;

namespace {

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t n) { return n <= cap ? cudaSuccess : allocate(n + n / 4 + 256); }
  // no headroom: for a buffer that must match another one's capacity
  cudaError_t reserve_exact(size_t n) { return n <= cap ? cudaSuccess : allocate(n); }
  cudaError_t allocate(size_t want) {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const {
    return reinterpret_cast<T*>(p);
  }
};

struct PinBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t n) {
    if (n <= cap) return cudaSuccess;
    size_t want = n + n / 4 + 256;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocDefault);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
  }
};

struct Slot {
  DevBuf data, off;
  uint32_t n = 0;
  uint64_t bytes = 0;
  // documents per producing tuple when the slot holds a generator's result (kinds repeat
  // with this period: StatefulSet, Service, VirtualService, ...); 1 for uploaded documents
  uint32_t period = 1;
  // nbrm_upload_async: the copy in flight on the upload stream, and the last generator
  // that read the slot (the next asynchronous copy into it waits for that one)
  cudaEvent_t uploaded = nullptr, last_read = nullptr;
  bool upload_pending = false, read_recorded = false;
  // structural index of the slot's documents (nbrm_slot_index), valid until the contents change
  DevBuf tape, tidx, tcursor;
  bool indexed = false;
  // views the producing generator left of the slot's documents (nbrm_plan.cuh), moved here by nbrm_result_to_slot
  DevBuf views;
  bool has_views = false;
};

// One download lane (nbrm_download_begin / _wait): the device buffers it took over
// from the result, its pinned host buffers, and the events that order the copy.
struct Lane {
  DevBuf data, off;
  DevBuf d_status, d_result;     // the lane's copy of the status / result words
  PinBuf h_out, h_off, h_status, h_result;
  cudaEvent_t ready = nullptr;   // compute stream: the result is complete
  cudaEvent_t done = nullptr;    // copy stream: the documents are in host memory
  cudaEvent_t small = nullptr;   // compute stream: status / result words are in host memory
  uint32_t n_in = 0, n_out = 0;
  uint64_t bytes = 0;
  int src_slot = -1;             // >= 0: the copy reads this input slot
  bool pending = false;
};

enum KClass { K_PLAN = 0, K_SCAN = 1, K_EMIT = 2, K_COUNT = 3 };

struct PendingEvent {
  int cls;
  int gen;
  cudaEvent_t a, b;
};

}  // namespace

struct nbrm_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t t0 = nullptr, t1 = nullptr;
  std::string err;

  // literal pool: fixed literals then the escaped environment strings
  DevBuf pool;
  EnvLits env;

  Slot slot[NBRM_SLOTS];
  DevBuf segs, nseg, lens, status, result, aux, out_off, scan_tmp, out, extra;
  DevBuf views;       // views of the last result's objects (generators that write them)
  bool res_views = false;
  size_t views_max = 0;  // view buffers rotate between the context and the slots: all grow to the largest asked for
  bool group_by_length = false;  // NBRM_OPT_GROUP_BY_LENGTH
  DevBuf group_buf, group_tmp;
  DevBuf gen_status;  // nbrm_nbc_reconcile: the generator stage's status words, kept across the compare
  uint64_t extra_bytes = 0;
  PinBuf h_out, h_off, h_status, h_result, h_total;
  Lane lane[NBRM_DL_LANES];
  cudaStream_t copy_stream = nullptr;
  cudaStream_t up_stream = nullptr;

  // last *_resident result
  uint32_t res_n_in = 0, res_n_out = 0, res_nobj = 1;
  uint64_t res_bytes = 0;
  bool res_valid = false;

  bool prof = false;
  std::vector<PendingEvent> pending;
  std::vector<cudaEvent_t> free_events;
  nbrm_profile acc;
  uint64_t launches = 0;
};

namespace {

int fail(nbrm_ctx* c, const char* what, cudaError_t e) {
  char buf[512];
  snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(e));
  c->err = buf;
  return -1;
}

int fail_msg(nbrm_ctx* c, const char* what) {
  c->err = what;
  return -2;
}

#define CK(call)                                   \
  do {                                             \
    cudaError_t e_ = (call);                       \
    if (e_ != cudaSuccess) return fail(c, #call, e_); \
  } while (0)

// A download lane that is still reading input slot `slot`: finish it before the
// slot's buffers are replaced.
int settle_slot_readers(nbrm_ctx* c, int slot) {
  for (auto& l : c->lane)
    if (l.pending && l.src_slot == slot) CK(cudaEventSynchronize(l.done));
  return 0;
}

// Work queued on the compute stream from here on sees the slot's asynchronous upload.
int order_after_upload(nbrm_ctx* c, Slot& s) {
  if (s.upload_pending) {
    CK(cudaStreamWaitEvent(c->stream, s.uploaded, 0));
    s.upload_pending = false;
  }
  return 0;
}

int lane_prepare(nbrm_ctx* c, int lane_id, const char* who) {
  if (lane_id < 0 || lane_id >= NBRM_DL_LANES) return fail_msg(c, "download lane out of range");
  Lane& l = c->lane[lane_id];
  if (l.pending) {
    c->err = std::string(who) + ": the lane's previous download was not waited for";
    return -2;
  }
  CK(cudaSetDevice(c->device));
  if (!c->copy_stream) CK(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
  if (!l.ready) {
    CK(cudaEventCreateWithFlags(&l.ready, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&l.done, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&l.small, cudaEventDisableTiming));
  }
  return 0;
}

cudaEvent_t get_event(nbrm_ctx* c) {
  if (!c->free_events.empty()) {
    cudaEvent_t e = c->free_events.back();
    c->free_events.pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}

struct KScope {  // brackets one kernel class with events when profiling is on
  nbrm_ctx* c;
  PendingEvent pe;
  bool on;
  KScope(nbrm_ctx* c_, int cls, uint32_t n_launches, int gen = 0) : c(c_), on(c_->prof) {
    c->launches += n_launches;
    if (!on) return;
    pe.cls = cls;
    pe.gen = gen & (NBRM_GENERATORS - 1);
    pe.a = get_event(c);
    pe.b = get_event(c);
    cudaEventRecord(pe.a, c->stream);
    if (cls == K_PLAN) c->acc.plan_launches += n_launches;
    if (cls == K_SCAN) c->acc.scan_launches += n_launches;
    if (cls == K_EMIT) c->acc.emit_launches += n_launches;
  }
  ~KScope() {
    if (!on) return;
    cudaEventRecord(pe.b, c->stream);
    c->pending.push_back(pe);
  }
};

int drain_profile(nbrm_ctx* c) {
  if (c->pending.empty()) return 0;
  CK(cudaStreamSynchronize(c->stream));
  for (auto& pe : c->pending) {
    float ms = 0;
    cudaEventElapsedTime(&ms, pe.a, pe.b);
    if (pe.cls == K_PLAN) {
      c->acc.plan_ms += ms;
      c->acc.plan_ms_gen[pe.gen & (NBRM_GENERATORS - 1)] += ms;
    }
    if (pe.cls == K_SCAN) c->acc.scan_ms += ms;
    if (pe.cls == K_EMIT) {
      c->acc.emit_ms += ms;
      c->acc.emit_ms_gen[pe.gen & (NBRM_GENERATORS - 1)] += ms;
    }
    c->free_events.push_back(pe.a);
    c->free_events.push_back(pe.b);
  }
  c->pending.clear();
  return 0;
}

int build_pool(nbrm_ctx* c, const nbrm_env* env) {
  std::string pool;
  build_pool_host(env, pool, c->env);
  CK(c->pool.reserve(pool.size()));
  CK(cudaMemcpyAsync(c->pool.p, pool.data(), pool.size(), cudaMemcpyHostToDevice, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

int check_docs(nbrm_ctx* c, const nbrm_docs* d) {
  if (!d || (d->n && (!d->data || !d->off))) return fail_msg(c, "nbrm_docs: null pointer");
  if (d->n > 0x7fffffffull / 4) return fail_msg(c, "nbrm_docs: too many documents in one batch");
  if (d->n) {
    if (d->off[d->n] - d->off[0] >= 0xffffffffull - d->n * AUX_PAD)
      return fail_msg(c, "nbrm_docs: batch holds 4 GiB or more; split it");
    for (uint64_t i = 0; i < d->n; ++i)
      if (d->off[i + 1] < d->off[i]) return fail_msg(c, "nbrm_docs: offsets must not decrease");
  }
  return 0;
}

}  // namespace

extern "C" {

int nbrm_ctx_create(int device, nbrm_ctx** out) {
  if (!out) return -2;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= 0) {
    fprintf(stderr, "libnbrm: no usable CUDA device (%s); there is no CPU fallback\n",
            e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return -1;
  }
  if (device < 0 || device >= ndev) return -2;
  nbrm_ctx* c = new nbrm_ctx();
  c->device = device;
  memset(&c->acc, 0, sizeof c->acc);
  if (cudaSetDevice(device) != cudaSuccess ||
      cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&c->t0) != cudaSuccess || cudaEventCreate(&c->t1) != cudaSuccess) {
    fprintf(stderr, "libnbrm: cannot set up device %d: %s\n", device,
            cudaGetErrorString(cudaGetLastError()));
    delete c;
    return -1;
  }
  cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
  for (auto& s : c->slot) {  // nbrm_upload_async orders itself against readers through these
    if (cudaEventCreateWithFlags(&s.uploaded, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s.last_read, cudaEventDisableTiming) != cudaSuccess) {
      fprintf(stderr, "libnbrm: cannot create events on device %d\n", device);
      nbrm_ctx_destroy(c);
      return -1;
    }
  }
  if (build_pool(c, nullptr) != 0) {
    fprintf(stderr, "libnbrm: %s\n", c->err.c_str());
    delete c;
    return -1;
  }
  *out = c;
  return 0;
}

void nbrm_ctx_destroy(nbrm_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  for (auto& pe : c->pending) {
    cudaEventDestroy(pe.a);
    cudaEventDestroy(pe.b);
  }
  for (auto e : c->free_events) cudaEventDestroy(e);
  if (c->up_stream) {
    cudaStreamSynchronize(c->up_stream);
    cudaStreamDestroy(c->up_stream);
  }
  for (auto& s : c->slot) {
    s.data.release();
    s.off.release();
    s.tape.release();
    s.tidx.release();
    s.tcursor.release();
    s.views.release();
    if (s.uploaded) cudaEventDestroy(s.uploaded);
    if (s.last_read) cudaEventDestroy(s.last_read);
  }
  DevBuf* bufs[] = {&c->pool, &c->segs, &c->nseg, &c->lens, &c->status,
                    &c->aux,  &c->out_off, &c->scan_tmp, &c->out, &c->extra, &c->result, &c->gen_status, &c->views};
  for (auto b : bufs) b->release();
  PinBuf* pins[] = {&c->h_out, &c->h_off, &c->h_status, &c->h_result, &c->h_total};
  for (auto b : pins) b->release();
  if (c->copy_stream) {
    cudaStreamSynchronize(c->copy_stream);
    cudaStreamDestroy(c->copy_stream);
  }
  for (auto& l : c->lane) {
    l.data.release();
    l.off.release();
    l.d_status.release();
    l.d_result.release();
    l.h_out.release();
    l.h_off.release();
    l.h_status.release();
    l.h_result.release();
    if (l.ready) cudaEventDestroy(l.ready);
    if (l.done) cudaEventDestroy(l.done);
    if (l.small) cudaEventDestroy(l.small);
  }
  cudaEventDestroy(c->t0);
  cudaEventDestroy(c->t1);
  cudaStreamDestroy(c->stream);
  delete c;
}

const char* nbrm_last_error(const nbrm_ctx* c) { return c ? c->err.c_str() : "null context"; }

int nbrm_ctx_set_option(nbrm_ctx* c, int option, int value) {
  if (!c) return -2;
  if (option == NBRM_OPT_GROUP_BY_LENGTH) {
    c->group_by_length = value != 0;
    return 0;
  }
  return fail_msg(c, "unknown option");
}

int nbrm_ctx_set_env(nbrm_ctx* c, const nbrm_env* env) {
  if (!c) return -2;
  CK(cudaSetDevice(c->device));
  return build_pool(c, env);
}

int nbrm_sync(nbrm_ctx* c) {
  if (!c) return -2;
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

void* nbrm_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) return nullptr;
  return p;
}

void nbrm_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

int nbrm_timer_start(nbrm_ctx* c) {
  if (!c) return -2;
  CK(cudaEventRecord(c->t0, c->stream));
  return 0;
}

int nbrm_timer_stop(nbrm_ctx* c, float* ms) {
  if (!c || !ms) return -2;
  CK(cudaEventRecord(c->t1, c->stream));
  CK(cudaEventSynchronize(c->t1));
  CK(cudaEventElapsedTime(ms, c->t0, c->t1));
  return 0;
}

int nbrm_profile_enable(nbrm_ctx* c, int on) {
  if (!c) return -2;
  if (drain_profile(c)) return -1;
  c->prof = on != 0;
  if (on) memset(&c->acc, 0, sizeof c->acc);
  return 0;
}

int nbrm_profile_read(nbrm_ctx* c, nbrm_profile* out) {
  if (!c || !out) return -2;
  if (drain_profile(c)) return -1;
  *out = c->acc;
  return 0;
}

uint64_t nbrm_launch_count(const nbrm_ctx* c) { return c ? c->launches : 0; }

int nbrm_upload(nbrm_ctx* c, int slot, const nbrm_docs* d) {
  if (!c) return -2;
  if (slot < 0 || slot >= NBRM_SLOTS) return fail_msg(c, "nbrm_upload: slot out of range");
  if (int r = check_docs(c, d)) return r;
  CK(cudaSetDevice(c->device));
  if (int r = settle_slot_readers(c, slot)) return r;
  Slot& s = c->slot[slot];
  if (int r = order_after_upload(c, s)) return r;
  s.indexed = false;
  s.has_views = false;
  s.n = (uint32_t)d->n;
  s.period = 1;
  uint64_t base = d->n ? d->off[0] : 0;
  if (base != 0) return fail_msg(c, "nbrm_docs: off[0] must be 0");
  s.bytes = d->n ? d->off[d->n] : 0;
  CK(s.data.reserve(s.bytes + 16));
  CK(s.off.reserve((d->n + 1) * sizeof(uint64_t)));
  if (s.bytes)
    CK(cudaMemcpyAsync(s.data.p, d->data, s.bytes, cudaMemcpyHostToDevice, c->stream));
  if (d->n) {
    CK(cudaMemcpyAsync(s.off.p, d->off, (d->n + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice,
                       c->stream));
  } else {
    CK(cudaMemsetAsync(s.off.p, 0, sizeof(uint64_t), c->stream));
  }
  c->res_valid = false;
  return 0;
}

int nbrm_upload_async(nbrm_ctx* c, int slot, const nbrm_docs* d) {
  if (!c) return -2;
  if (slot < 0 || slot >= NBRM_SLOTS) return fail_msg(c, "nbrm_upload_async: slot out of range");
  if (int r = check_docs(c, d)) return r;
  if (d->n && d->off[0] != 0) return fail_msg(c, "nbrm_docs: off[0] must be 0");
  CK(cudaSetDevice(c->device));
  if (int r = settle_slot_readers(c, slot)) return r;
  Slot& s = c->slot[slot];
  if (!c->up_stream) CK(cudaStreamCreateWithFlags(&c->up_stream, cudaStreamNonBlocking));
  // generators already launched may still read the slot's old contents
  if (s.read_recorded) CK(cudaStreamWaitEvent(c->up_stream, s.last_read, 0));
  const uint64_t bytes = d->n ? d->off[d->n] : 0;
  if (bytes + 16 > s.data.cap || (d->n + 1) * sizeof(uint64_t) > s.off.cap) {
    // growing frees the old buffer: nothing may be in flight on it
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaStreamSynchronize(c->up_stream));
  }
  CK(s.data.reserve(bytes + 16));
  CK(s.off.reserve((d->n + 1) * sizeof(uint64_t)));
  s.indexed = false;
  s.has_views = false;
  s.n = (uint32_t)d->n;
  s.period = 1;
  s.bytes = bytes;
  if (bytes) CK(cudaMemcpyAsync(s.data.p, d->data, bytes, cudaMemcpyHostToDevice, c->up_stream));
  if (d->n) {
    CK(cudaMemcpyAsync(s.off.p, d->off, (d->n + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice,
                       c->up_stream));
  } else {
    CK(cudaMemsetAsync(s.off.p, 0, sizeof(uint64_t), c->up_stream));
  }
  CK(cudaEventRecord(s.uploaded, c->up_stream));
  s.upload_pending = true;
  return 0;
}

}  // extern "C"

namespace {

// plan -> scan -> (read total) -> emit, shared by every generator.
// map[k] = the context slot that feeds generator input k, or -1 (all used slots
// must hold the same number of documents; tuple i is document i of each).
int run_generator(nbrm_ctx* c, int gen, const int (&map)[3], uint32_t flags, const void* extra,
                  uint64_t now, uint32_t arg0 = 0, uint32_t arg1 = 0) {
  CK(cudaSetDevice(c->device));
  uint32_t nobj = 0, cap = 0;
  gen_shape(gen, &nobj, &cap);
  uint32_t n = 0;
  uint64_t in_bytes = 0;
  bool first = true;
  PlanIn in;
  memset(&in, 0, sizeof in);
  for (int k = 0; k < 3; ++k) {
    if (map[k] < 0) continue;
    Slot& s = c->slot[map[k]];
    if (!s.off.p) return fail_msg(c, "an input slot this entry point reads was never uploaded");
    if (int r = order_after_upload(c, s)) return r;
    const bool shared = k == 1 && gen_shared1(gen);
    if (shared) {
      in.shared_n = s.n;
    } else {
      if (first) n = s.n;
      else if (s.n != n) return fail_msg(c, "input slots hold different numbers of documents");
      first = false;
    }
    in.slot[k].data = s.data.as<char>();
    in.slot[k].off = s.off.as<uint64_t>();
    if (s.indexed) {  // per-tuple slots and the shared arena alike
      in.slot[k].tape = s.tape.p;
      in.slot[k].tidx = s.tidx.as<uint2>();
    }
    if (s.has_views && !shared) in.slot[k].views = s.views.as<ViewEnt>();
    in_bytes += s.bytes;
  }
  if (in_bytes + (uint64_t)n * AUX_PAD >= 0xffffffffull)
    return fail_msg(c, "inputs of this batch total 4 GiB or more; split it");
  in.n = n;
  in.flags = flags;
  in.env = c->env;
  in.extra = extra;
  in.now = now;
  in.arg0 = arg0;
  in.arg1 = arg1;
  in.period = map[0] >= 0 ? c->slot[map[0]].period : 1;
  in.perm = nullptr;
  if (c->group_by_length && map[0] >= 0 && in.period == 1 && n > 4096) {
    // threads take the tuples in the order of their first document's length (nbrm_group.cu)
    const size_t tb = group_tmp_bytes(n);
    CK(c->group_buf.reserve((size_t)n * 16 + 16));
    CK(c->group_tmp.reserve(tb));
    in.perm = launch_group_by_length(c->slot[map[0]].off.as<uint64_t>(), n, c->group_buf.as<uint32_t>(),
                                     c->group_tmp.p, tb, c->stream);
    c->launches += 1;  // the key kernel; the sort's kernels are the library's
    CK(cudaGetLastError());
  }
  const uint32_t n_out = n * nobj;
  CK(c->segs.reserve((size_t)n_out * cap * sizeof(Seg) + 16));
  CK(c->nseg.reserve((size_t)n_out * 4 + 16));
  CK(c->lens.reserve((size_t)n_out * 4 + 16));
  CK(c->status.reserve((size_t)n * 4 + 16));
  CK(c->result.reserve((size_t)n * 4 + 16));
  CK(c->aux.reserve(in_bytes + (uint64_t)n * AUX_PAD + 16));
  CK(c->out_off.reserve(((size_t)n_out + 1) * 8));
  CK(c->scan_tmp.reserve(scan_tmp_bytes(n_out)));
  CK(c->h_total.reserve(8));
  PlanOut po;
  po.segs = c->segs.as<Seg>();
  po.nseg = c->nseg.as<uint32_t>();
  po.lens = c->lens.as<uint32_t>();
  po.status = c->status.as<uint32_t>();
  po.result = c->result.as<uint32_t>();
  po.aux = c->aux.as<char>();
  po.cap = cap;
  po.views = nullptr;
  c->res_views = false;
  if (gen_views(gen)) {
    const size_t want = (size_t)n_out * VIEW_FIELDS * sizeof(ViewEnt) + 16;
    if (want > c->views_max) c->views_max = want + want / 4;
    CK(c->views.reserve_exact(c->views_max));
    po.views = c->views.as<ViewEnt>();
    // planners store only the fields an object has; everything else reads as an empty span
    CK(cudaMemsetAsync(c->views.p, 0, (size_t)n_out * VIEW_FIELDS * sizeof(ViewEnt), c->stream));
    c->res_views = true;
  }
  {
    KScope k(c, K_PLAN, n ? 1 : 0, gen);
    launch_plan(gen, in, po, in_bytes, c->stream);
  }
  CK(cudaGetLastError());
  {
    KScope k(c, K_SCAN, n_out ? 3 : 0);
    launch_scan(po.lens, n_out, c->scan_tmp.as<uint64_t>(), c->out_off.as<uint64_t>(), c->stream);
  }
  CK(cudaGetLastError());
  // the SM stores the total into pinned host memory (see launch_store_u64)
  launch_store_u64(c->out_off.as<uint64_t>() + n_out, reinterpret_cast<uint64_t*>(c->h_total.p), c->stream);
  c->launches += 1;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(c->stream));
  uint64_t total = *reinterpret_cast<uint64_t*>(c->h_total.p);
  if (total >= 0xffffffffull) return fail_msg(c, "output of this batch is 4 GiB or more; split it");
  CK(c->out.reserve(total + 16));
  SrcBases b;
  b.p[SRC_IN0] = map[0] >= 0 ? c->slot[map[0]].data.as<char>() : nullptr;
  b.p[SRC_IN1] = map[1] >= 0 ? c->slot[map[1]].data.as<char>() : nullptr;
  b.p[SRC_IN2] = map[2] >= 0 ? c->slot[map[2]].data.as<char>() : nullptr;
  b.p[SRC_POOL] = c->pool.as<char>();
  b.p[SRC_AUX] = c->aux.as<char>();
  {
    KScope k(c, K_EMIT, n_out ? 1 : 0, gen);
    launch_emit(po.segs, po.nseg, po.lens, c->out_off.as<uint64_t>(), n_out, cap, nobj, b,
                c->out.as<char>(), c->sm_count, c->stream);
  }
  CK(cudaGetLastError());
  if (c->prof) {
    c->acc.emit_bytes_read += total;
    c->acc.emit_bytes_written += total;
    c->acc.plan_bytes_read += in_bytes;
    c->acc.plan_bytes_gen[gen & (NBRM_GENERATORS - 1)] += in_bytes;
    c->acc.emit_bytes_gen[gen & (NBRM_GENERATORS - 1)] += 2 * total;
    c->acc.runs_gen[gen & (NBRM_GENERATORS - 1)] += 1;
  }
  for (int k = 0; k < 3; ++k) {  // for nbrm_upload_async: this generator read these slots
    if (map[k] < 0) continue;
    Slot& s = c->slot[map[k]];
    CK(cudaEventRecord(s.last_read, c->stream));
    s.read_recorded = true;
  }
  c->res_n_in = n;
  c->res_n_out = n_out;
  c->res_nobj = nobj;
  c->res_bytes = total;
  c->res_valid = true;
  return 0;
}

}  // namespace

extern "C" {

int nbrm_download(nbrm_ctx* c, nbrm_out* out) {
  if (!c || !out) return -2;
  if (!c->res_valid) return fail_msg(c, "nbrm_download: no result on this context");
  CK(cudaSetDevice(c->device));
  CK(c->h_out.reserve(c->res_bytes + 16));
  CK(c->h_off.reserve(((size_t)c->res_n_out + 1) * 8));
  CK(c->h_status.reserve((size_t)c->res_n_in * 4 + 16));
  CK(c->h_result.reserve((size_t)c->res_n_in * 4 + 16));
  if (c->res_bytes)
    CK(cudaMemcpyAsync(c->h_out.p, c->out.p, c->res_bytes, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(c->h_off.p, c->out_off.p, ((size_t)c->res_n_out + 1) * 8,
                     cudaMemcpyDeviceToHost, c->stream));
  if (c->res_n_in) {
    CK(cudaMemcpyAsync(c->h_status.p, c->status.p, (size_t)c->res_n_in * 4, cudaMemcpyDeviceToHost,
                       c->stream));
    CK(cudaMemcpyAsync(c->h_result.p, c->result.p, (size_t)c->res_n_in * 4, cudaMemcpyDeviceToHost,
                       c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  out->data = reinterpret_cast<const char*>(c->h_out.p);
  out->off = reinterpret_cast<const uint64_t*>(c->h_off.p);
  out->n = c->res_n_out;
  out->status = reinterpret_cast<const uint32_t*>(c->h_status.p);
  out->n_status = c->res_n_in;
  out->result = reinterpret_cast<const uint32_t*>(c->h_result.p);
  const uint32_t* st = out->status;
  for (uint32_t i = 0; i < c->res_n_in; ++i)
    if (st[i]) return 1;
  return 0;
}

int nbrm_upload_bytes(nbrm_ctx* c, const void* host, uint64_t bytes) {
  if (!c || (bytes && !host)) return -2;
  CK(cudaSetDevice(c->device));
  CK(c->extra.reserve(bytes + 16));
  if (bytes) CK(cudaMemcpyAsync(c->extra.p, host, bytes, cudaMemcpyHostToDevice, c->stream));
  c->extra_bytes = bytes;
  return 0;
}

int nbrm_nbc_generate_resident(nbrm_ctx* c, uint32_t flags) {
  if (!c) return -2;
  return run_generator(c, GEN_ID_NBC, {0, -1, -1}, flags, nullptr, 0);
}

int nbrm_nbc_generate(nbrm_ctx* c, const nbrm_docs* notebooks, uint32_t flags, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, notebooks)) return r;
  if (int r = nbrm_nbc_generate_resident(c, flags)) return r;
  return nbrm_download(c, out);
}

int nbrm_odh_generate_resident(nbrm_ctx* c, uint32_t flags) {
  if (!c) return -2;
  const void* seeds = nullptr;
  if (c->extra_bytes) {
    if (c->extra_bytes != (uint64_t)c->slot[0].n * 16)
      return fail_msg(c, "nbrm_odh_generate: seeds must hold 16 bytes per Notebook");
    seeds = c->extra.p;
  }
  return run_generator(c, GEN_ID_ODH, {0, -1, -1}, flags, seeds, 0);
}

int nbrm_odh_generate(nbrm_ctx* c, const nbrm_docs* notebooks, const uint8_t* seeds16,
                      uint32_t flags, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, notebooks)) return r;
  if (int r = nbrm_upload_bytes(c, seeds16, seeds16 ? notebooks->n * 16 : 0)) return r;
  if (int r = nbrm_odh_generate_resident(c, flags)) return r;
  return nbrm_download(c, out);
}

int nbrm_result_to_slot(nbrm_ctx* c, int slot) {
  if (!c) return -2;
  if (slot < 0 || slot >= NBRM_SLOTS) return fail_msg(c, "nbrm_result_to_slot: slot out of range");
  if (!c->res_valid) return fail_msg(c, "nbrm_result_to_slot: no result on this context");
  if (int r = settle_slot_readers(c, slot)) return r;
  Slot& s = c->slot[slot];
  if (int r = order_after_upload(c, s)) return r;
  std::swap(s.data, c->out);
  std::swap(s.off, c->out_off);
  std::swap(s.views, c->views);
  s.has_views = c->res_views;
  c->res_views = false;
  s.indexed = false;
  // the buffer now in the slot is still being written by the emit kernel, and the one that left
  // it may still be read: an asynchronous upload into this slot waits for the compute stream
  CK(cudaEventRecord(s.last_read, c->stream));
  s.read_recorded = true;
  s.n = c->res_n_out;
  s.period = c->res_nobj;
  s.bytes = c->res_bytes;
  c->res_valid = false;
  return 0;
}

int nbrm_copy_fields_resident(nbrm_ctx* c) {
  if (!c) return -2;
  return run_generator(c, GEN_ID_COPY, {0, 1, -1}, 0, nullptr, 0);
}

int nbrm_copy_fields(nbrm_ctx* c, const nbrm_docs* desired, const nbrm_docs* found, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, desired)) return r;
  if (int r = nbrm_upload(c, 1, found)) return r;
  if (int r = nbrm_copy_fields_resident(c)) return r;
  return nbrm_download(c, out);
}

int nbrm_nbc_reconcile_resident(nbrm_ctx* c, uint32_t flags) {
  if (!c) return -2;
  if (int r = run_generator(c, GEN_ID_NBC, {0, -1, -1}, flags, nullptr, 0)) return r;
  // the compare stage overwrites the status words: keep the generators' (E_NO_CONTAINERS,
  // E_PORTS_EMPTY, E_PARSE ... where the reference panics before it writes anything)
  const uint32_t n_nb = c->res_n_in;
  CK(c->gen_status.reserve((size_t)n_nb * 4 + 16));
  launch_copy_u32(c->status.as<uint32_t>(), c->gen_status.as<uint32_t>(), n_nb, c->stream);
  c->launches += n_nb ? 1 : 0;
  if (int r = nbrm_result_to_slot(c, 2)) return r;  // desired objects become input slot 2
  if (int r = run_generator(c, GEN_ID_COPY, {2, 1, -1}, 0, nullptr, 0)) return r;
  launch_merge_status(c->gen_status.as<uint32_t>(), c->status.as<uint32_t>(), c->res_n_in, 3, c->stream);
  c->launches += c->res_n_in ? 1 : 0;
  CK(cudaGetLastError());
  return 0;
}

int nbrm_nbc_reconcile(nbrm_ctx* c, const nbrm_docs* notebooks, const nbrm_docs* found,
                       uint32_t flags, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, notebooks)) return r;
  if (int r = nbrm_upload(c, 1, found)) return r;
  if (int r = nbrm_nbc_reconcile_resident(c, flags)) return r;
  return nbrm_download(c, out);
}

int nbrm_notebook_status_resident(nbrm_ctx* c, uint64_t now_unix) {
  if (!c) return -2;
  return run_generator(c, GEN_ID_STATUS, {0, 1, 2}, 0, nullptr, now_unix);
}

int nbrm_notebook_status(nbrm_ctx* c, const nbrm_docs* notebooks, const nbrm_docs* statefulsets,
                         const nbrm_docs* pods, uint64_t now_unix, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, notebooks)) return r;
  if (int r = nbrm_upload(c, 1, statefulsets)) return r;
  if (int r = nbrm_upload(c, 2, pods)) return r;
  if (int r = nbrm_notebook_status_resident(c, now_unix)) return r;
  return nbrm_download(c, out);
}

int nbrm_culling_pass_resident(nbrm_ctx* c, uint64_t now_unix, int32_t cull_idle_minutes,
                               int32_t check_period_minutes) {
  if (!c) return -2;
  const void* pod_exists = nullptr;
  if (c->extra_bytes) {
    if (c->extra_bytes != (uint64_t)c->slot[0].n)
      return fail_msg(c, "nbrm_culling_pass: pod_exists must hold one byte per Notebook");
    pod_exists = c->extra.p;
  }
  return run_generator(c, GEN_ID_CULL, {0, 1, 2}, 0, pod_exists, now_unix, (uint32_t)cull_idle_minutes,
                       (uint32_t)check_period_minutes);
}

int nbrm_culling_pass(nbrm_ctx* c, const nbrm_docs* notebooks, const nbrm_docs* kernels,
                      const nbrm_docs* terminals, const uint8_t* pod_exists, uint64_t now_unix,
                      int32_t cull_idle_minutes, int32_t check_period_minutes, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, notebooks)) return r;
  if (int r = nbrm_upload(c, 1, kernels)) return r;
  if (int r = nbrm_upload(c, 2, terminals)) return r;
  if (int r = nbrm_upload_bytes(c, pod_exists, pod_exists ? notebooks->n : 0)) return r;
  if (int r = nbrm_culling_pass_resident(c, now_unix, cull_idle_minutes, check_period_minutes)) return r;
  return nbrm_download(c, out);
}

int nbrm_odh_mutate_resident(nbrm_ctx* c, uint32_t flags, int is_update) {
  if (!c) return -2;
  if (!is_update) return run_generator(c, GEN_ID_MUTATE, {1, -1, -1}, flags | 1u, nullptr, 0);
  if (int r = run_generator(c, GEN_ID_MUTATE, {1, -1, -1}, flags & ~1u, nullptr, 0)) return r;
  if (int r = nbrm_result_to_slot(c, 0)) return r;  // the mutated Notebooks become slot 0
  return run_generator(c, GEN_ID_GUARD, {0, 1, 2}, 0, nullptr, 0);
}

int nbrm_odh_mutate(nbrm_ctx* c, const nbrm_docs* notebooks, const nbrm_docs* old_notebooks,
                    uint32_t flags, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 1, notebooks)) return r;
  if (old_notebooks)
    if (int r = nbrm_upload(c, 2, old_notebooks)) return r;
  if (int r = nbrm_odh_mutate_resident(c, flags, old_notebooks != nullptr)) return r;
  return nbrm_download(c, out);
}

int nbrm_poddefault_mutate_resident(nbrm_ctx* c) {
  if (!c) return -2;
  // every Pod of the batch consults every PodDefault of its namespace several times: the list is
  // indexed once (it usually stays resident across batches, like the reference's List would
  // come from a cache) and the planner hops through the index instead of re-reading its text
  if (c->slot[1].off.p && !c->slot[1].indexed)
    if (int r = nbrm_slot_index(c, 1)) return r;
  const void* ranges = nullptr;
  if (c->extra_bytes) {
    if (c->extra_bytes != (uint64_t)c->slot[0].n * 8)
      return fail_msg(c, "nbrm_poddefault_mutate: ranges must hold two uint32 per Pod");
    ranges = c->extra.p;
  }
  return run_generator(c, GEN_ID_PODDEFAULT, {0, 1, -1}, 0, ranges, 0);
}

int nbrm_poddefault_mutate(nbrm_ctx* c, const nbrm_docs* pods, const nbrm_docs* poddefaults,
                           const uint32_t* ranges, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, pods)) return r;
  if (int r = nbrm_upload(c, 1, poddefaults)) return r;
  if (int r = nbrm_upload_bytes(c, ranges, ranges ? pods->n * 8 : 0)) return r;
  if (int r = nbrm_poddefault_mutate_resident(c)) return r;
  return nbrm_download(c, out);
}

int nbrm_run(nbrm_ctx* c, int generator, int slot_a, int slot_b, int slot_c, uint32_t flags,
             uint64_t now_unix, uint32_t arg0, uint32_t arg1, int use_side_input) {
  if (!c) return -2;
  uint32_t nobj = 0, cap = 0;
  gen_shape(generator, &nobj, &cap);
  if (!nobj) return fail_msg(c, "nbrm_run: unknown generator");
  const int map[3] = {slot_a, slot_b, slot_c};
  for (int k = 0; k < 3; ++k)
    if (map[k] >= NBRM_SLOTS) return fail_msg(c, "nbrm_run: slot out of range");
  if (slot_a < 0) return fail_msg(c, "nbrm_run: the first input slot is required");
  const void* side = nullptr;
  if (use_side_input && c->extra_bytes) {
    // the side input is read at a generator-specific stride: check it against this batch
    const uint64_t n = c->slot[slot_a].n;
    uint64_t want = 0;
    switch (generator) {
      case GEN_ID_ODH: want = n * 16; break;        // 16 seed bytes per Notebook
      case GEN_ID_CULL: want = n; break;            // pod_exists, one byte per Notebook
      case GEN_ID_PODDEFAULT: want = n * 8; break;  // [begin, end) per Pod
      default: return fail_msg(c, "nbrm_run: this generator takes no side input");
    }
    if (c->extra_bytes != want)
      return fail_msg(c, "nbrm_run: the side input uploaded with nbrm_upload_bytes does not fit this batch");
    side = c->extra.p;
  }
  // the bytes stay until nbrm_upload_bytes replaces them (a resident batch is planned many times)
  return run_generator(c, generator, map, flags, side, now_unix, arg0, arg1);
}

int nbrm_download_slot(nbrm_ctx* c, int slot, nbrm_out* out) {
  if (!c || !out) return -2;
  if (slot < 0 || slot >= NBRM_SLOTS) return fail_msg(c, "nbrm_download_slot: slot out of range");
  Slot& s = c->slot[slot];
  if (!s.off.p) return fail_msg(c, "nbrm_download_slot: the slot is empty");
  CK(cudaSetDevice(c->device));
  if (int r = order_after_upload(c, s)) return r;
  CK(c->h_out.reserve(s.bytes + 16));
  CK(c->h_off.reserve(((size_t)s.n + 1) * 8));
  if (s.bytes) CK(cudaMemcpyAsync(c->h_out.p, s.data.p, s.bytes, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(c->h_off.p, s.off.p, ((size_t)s.n + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  out->data = reinterpret_cast<const char*>(c->h_out.p);
  out->off = reinterpret_cast<const uint64_t*>(c->h_off.p);
  out->n = s.n;
  out->status = nullptr;
  out->n_status = 0;
  out->result = nullptr;
  return 0;
}

int nbrm_download_begin(nbrm_ctx* c, int lane_id) {
  if (!c) return -2;
  if (!c->res_valid) return fail_msg(c, "nbrm_download_begin: no result on this context");
  if (int r = lane_prepare(c, lane_id, "nbrm_download_begin")) return r;
  Lane& l = c->lane[lane_id];
  l.n_in = c->res_n_in;
  l.n_out = c->res_n_out;
  l.bytes = c->res_bytes;
  l.src_slot = -1;
  CK(l.h_out.reserve(l.bytes + 16));
  CK(l.h_off.reserve(((size_t)l.n_out + 1) * 8));
  CK(l.h_status.reserve((size_t)l.n_in * 4 + 16));
  CK(l.h_result.reserve((size_t)l.n_in * 4 + 16));
  // the buffers that come back to the context must be able to hold what it holds now,
  // so that after the first pass no generator has to reallocate its output
  CK(l.data.reserve_exact(c->out.cap));
  CK(l.off.reserve_exact(c->out_off.cap));
  // status / result words: the next generator overwrites them, so the lane takes a copy
  // on the compute stream (a kernel: a device-to-host copy here would queue behind another
  // lane's transfer in the copy engine and hold up the generators that follow)
  CK(l.d_status.reserve((size_t)l.n_in * 4 + 16));
  CK(l.d_result.reserve((size_t)l.n_in * 4 + 16));
  if (l.n_in) {
    launch_copy_u32(c->status.as<uint32_t>(), l.d_status.as<uint32_t>(), l.n_in, c->stream);
    launch_copy_u32(c->result.as<uint32_t>(), l.d_result.as<uint32_t>(), l.n_in, c->stream);
    c->launches += 2;
    CK(cudaGetLastError());
  }
  CK(cudaEventRecord(l.small, c->stream));
  CK(cudaEventRecord(l.ready, c->stream));
  std::swap(l.data, c->out);
  std::swap(l.off, c->out_off);
  c->res_valid = false;
  CK(cudaStreamWaitEvent(c->copy_stream, l.ready, 0));
  if (l.bytes) CK(cudaMemcpyAsync(l.h_out.p, l.data.p, l.bytes, cudaMemcpyDeviceToHost, c->copy_stream));
  CK(cudaMemcpyAsync(l.h_off.p, l.off.p, ((size_t)l.n_out + 1) * 8, cudaMemcpyDeviceToHost, c->copy_stream));
  if (l.n_in) {
    CK(cudaMemcpyAsync(l.h_status.p, l.d_status.p, (size_t)l.n_in * 4, cudaMemcpyDeviceToHost, c->copy_stream));
    CK(cudaMemcpyAsync(l.h_result.p, l.d_result.p, (size_t)l.n_in * 4, cudaMemcpyDeviceToHost, c->copy_stream));
  }
  CK(cudaEventRecord(l.done, c->copy_stream));
  l.pending = true;
  return 0;
}

int nbrm_download_slot_begin(nbrm_ctx* c, int slot, int lane_id) {
  if (!c) return -2;
  if (slot < 0 || slot >= NBRM_SLOTS) return fail_msg(c, "nbrm_download_slot_begin: slot out of range");
  Slot& s = c->slot[slot];
  if (!s.off.p) return fail_msg(c, "nbrm_download_slot_begin: the slot is empty");
  if (int r = lane_prepare(c, lane_id, "nbrm_download_slot_begin")) return r;
  if (int r = order_after_upload(c, s)) return r;
  Lane& l = c->lane[lane_id];
  l.n_in = 0;
  l.n_out = s.n;
  l.bytes = s.bytes;
  l.src_slot = slot;
  CK(l.h_out.reserve(l.bytes + 16));
  CK(l.h_off.reserve(((size_t)l.n_out + 1) * 8));
  CK(cudaEventRecord(l.small, c->stream));
  CK(cudaEventRecord(l.ready, c->stream));
  CK(cudaStreamWaitEvent(c->copy_stream, l.ready, 0));
  if (l.bytes) CK(cudaMemcpyAsync(l.h_out.p, s.data.p, l.bytes, cudaMemcpyDeviceToHost, c->copy_stream));
  CK(cudaMemcpyAsync(l.h_off.p, s.off.p, ((size_t)l.n_out + 1) * 8, cudaMemcpyDeviceToHost, c->copy_stream));
  CK(cudaEventRecord(l.done, c->copy_stream));
  l.pending = true;
  return 0;
}

int nbrm_download_wait(nbrm_ctx* c, int lane_id, nbrm_out* out) {
  if (!c || !out) return -2;
  if (lane_id < 0 || lane_id >= NBRM_DL_LANES) return fail_msg(c, "download lane out of range");
  Lane& l = c->lane[lane_id];
  if (!l.pending) return fail_msg(c, "nbrm_download_wait: nothing was begun on this lane");
  CK(cudaSetDevice(c->device));
  CK(cudaEventSynchronize(l.small));
  CK(cudaEventSynchronize(l.done));
  l.pending = false;
  out->data = reinterpret_cast<const char*>(l.h_out.p);
  out->off = reinterpret_cast<const uint64_t*>(l.h_off.p);
  out->n = l.n_out;
  out->status = l.n_in ? reinterpret_cast<const uint32_t*>(l.h_status.p) : nullptr;
  out->n_status = l.n_in;
  out->result = l.n_in ? reinterpret_cast<const uint32_t*>(l.h_result.p) : nullptr;
  for (uint32_t i = 0; i < l.n_in; ++i)
    if (out->status[i]) return 1;
  return 0;
}

// ---- wire formats (nbrm_patch.cuh)
static int status_to_extra(nbrm_ctx* c, const uint32_t* keep) {
  // the status words of the stage that made the patches become the side input of the
  // envelope stage; `keep` (earlier stage, may be null) wins where it is not 0
  const uint32_t n = c->res_n_in;
  if (keep) {
    launch_merge_status(keep, c->status.as<uint32_t>(), n, 1, c->stream);
    c->launches += n ? 1 : 0;
  }
  CK(c->extra.reserve((size_t)n * 4 + 16));
  launch_copy_u32(c->status.as<uint32_t>(), c->extra.as<uint32_t>(), n, c->stream);
  c->launches += n ? 1 : 0;
  c->extra_bytes = (uint64_t)n * 4;
  CK(cudaGetLastError());
  return 0;
}

int nbrm_json_diff(nbrm_ctx* c, const nbrm_docs* originals, const nbrm_docs* mutated, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, originals)) return r;
  if (int r = nbrm_upload(c, 1, mutated)) return r;
  if (int r = run_generator(c, GEN_ID_DIFF, {0, 1, -1}, 0, nullptr, 0)) return r;
  return nbrm_download(c, out);
}

int nbrm_notebook_webhook_review(nbrm_ctx* c, const nbrm_docs* reviews, uint32_t flags, int is_update,
                                 nbrm_out* out) {
  if (!c || !out) return -2;
  const int S_REVIEWS = 9, S_MUTATED = 10, S_PATCH = 11;
  if (int r = nbrm_upload(c, S_REVIEWS, reviews)) return r;
  if (int r = run_generator(c, GEN_ID_REVIEW_OPEN, {S_REVIEWS, -1, -1}, 0, nullptr, 0)) return r;
  if (int r = nbrm_result_to_slot(c, 1)) return r;  // request.object: what the mutation reads
  if (is_update) {
    if (int r = run_generator(c, GEN_ID_REVIEW_OPEN, {S_REVIEWS, -1, -1}, 2u /* REVIEW_OLD */, nullptr, 0)) return r;
    if (int r = nbrm_result_to_slot(c, 2)) return r;
  }
  if (int r = nbrm_odh_mutate_resident(c, flags, is_update)) return r;
  const uint32_t n = c->res_n_in;
  CK(c->gen_status.reserve((size_t)n * 4 + 16));
  launch_copy_u32(c->status.as<uint32_t>(), c->gen_status.as<uint32_t>(), n, c->stream);
  c->launches += n ? 1 : 0;
  if (int r = nbrm_result_to_slot(c, S_MUTATED)) return r;
  // admission.PatchResponseFromRaw(req.Object.Raw, marshaledNotebook) (notebook_webhook.go:301)
  if (int r = run_generator(c, GEN_ID_DIFF, {1, S_MUTATED, -1}, 0, nullptr, 0)) return r;
  if (int r = status_to_extra(c, c->gen_status.as<uint32_t>())) return r;
  if (int r = nbrm_result_to_slot(c, S_PATCH)) return r;
  if (int r = run_generator(c, GEN_ID_REVIEW_CLOSE, {S_REVIEWS, S_PATCH, -1}, 1u /* REVIEW_CTRL_RUNTIME */,
                            c->extra.p, 0))
    return r;
  c->extra_bytes = 0;
  return nbrm_download(c, out);
}

int nbrm_poddefault_review(nbrm_ctx* c, const nbrm_docs* reviews, const nbrm_docs* poddefaults,
                           const uint32_t* ranges, nbrm_out* out) {
  if (!c || !out) return -2;
  const int S_REVIEWS = 9, S_PATCH = 11;
  if (int r = nbrm_upload(c, S_REVIEWS, reviews)) return r;
  if (int r = nbrm_upload(c, 1, poddefaults)) return r;
  if (int r = run_generator(c, GEN_ID_REVIEW_OPEN, {S_REVIEWS, -1, -1}, 0, nullptr, 0)) return r;
  if (int r = nbrm_result_to_slot(c, 0)) return r;  // the Pods, namespace filled from the request
  if (int r = nbrm_upload_bytes(c, ranges, ranges ? reviews->n * 8 : 0)) return r;
  if (int r = nbrm_poddefault_mutate_resident(c)) return r;
  if (int r = status_to_extra(c, nullptr)) return r;
  if (int r = nbrm_result_to_slot(c, S_PATCH)) return r;
  // serve: uid echoed, patch base64, no response at all for anything but core/v1 pods (main.go:708-749)
  if (int r = run_generator(c, GEN_ID_REVIEW_CLOSE, {S_REVIEWS, S_PATCH, -1}, 4u /* REVIEW_PODS_ONLY */, c->extra.p, 0))
    return r;
  c->extra_bytes = 0;
  return nbrm_download(c, out);
}

int nbrm_notebook_convert(nbrm_ctx* c, const nbrm_docs* notebooks, int to_version, uint32_t flags, nbrm_out* out) {
  if (!c || !out) return -2;
  if (to_version < 0 || to_version > 2) return fail_msg(c, "nbrm_notebook_convert: to_version is 0 (v1beta1), 1 (v1) or 2 (v1alpha1)");
  if (int r = nbrm_upload(c, 0, notebooks)) return r;
  if (int r = run_generator(c, GEN_ID_CONVERT, {0, -1, -1}, flags, nullptr, 0, (uint32_t)to_version, 0)) return r;
  return nbrm_download(c, out);
}

int nbrm_slot_index(nbrm_ctx* c, int slot) {
  if (!c) return -2;
  if (slot < 0 || slot >= NBRM_SLOTS) return fail_msg(c, "nbrm_slot_index: slot out of range");
  Slot& s = c->slot[slot];
  if (!s.off.p) return fail_msg(c, "nbrm_slot_index: the slot is empty");
  CK(cudaSetDevice(c->device));
  if (int r = order_after_upload(c, s)) return r;
  if (s.indexed || !s.n) return 0;
  // API objects have one structural character per 8 to 11 bytes of compact JSON; room for
  // one per 5 bytes.  Documents that find the arena full simply stay without an index.
  const uint64_t want = s.bytes / 5 + (uint64_t)s.n * 8 + 64;
  const uint32_t arena_cap = want > 0xfffffff0ull ? 0xfffffff0u : (uint32_t)want;
  CK(s.tape.reserve((size_t)arena_cap * 8));
  CK(s.tidx.reserve((size_t)s.n * 8 + 16));
  CK(s.tcursor.reserve(16));
  DevDocs docs;
  memset(&docs, 0, sizeof docs);
  docs.data = s.data.as<char>();
  docs.off = s.off.as<uint64_t>();
  launch_index_slot(docs, s.n, s.tape.p, arena_cap, s.tcursor.as<uint32_t>(), s.tidx.p, c->stream);
  c->launches += 1;
  CK(cudaGetLastError());
  s.indexed = true;
  return 0;
}

int nbrm_tape_dump(nbrm_ctx* c, int slot, uint32_t cap, uint64_t* entries, uint32_t* counts, uint32_t* flags) {
  if (!c || !entries || !counts || !flags) return -2;
  if (slot < 0 || slot >= NBRM_SLOTS) return fail_msg(c, "nbrm_tape_dump: slot out of range");
  if (cap == 0 || cap > 24000) return fail_msg(c, "nbrm_tape_dump: cap must be 1..24000 entries");
  Slot& s = c->slot[slot];
  if (!s.off.p) return fail_msg(c, "nbrm_tape_dump: the slot is empty");
  CK(cudaSetDevice(c->device));
  if (int r = order_after_upload(c, s)) return r;
  if (!s.n) return 0;
  DevBuf d_ent, d_cnt, d_flg;
  cudaError_t e = d_ent.reserve_exact((size_t)s.n * cap * 8);
  if (e == cudaSuccess) e = d_cnt.reserve_exact((size_t)s.n * 4);
  if (e == cudaSuccess) e = d_flg.reserve_exact((size_t)s.n * 4);
  if (e == cudaSuccess) {
    DevDocs docs;
    memset(&docs, 0, sizeof docs);
    docs.data = s.data.as<char>();
    docs.off = s.off.as<uint64_t>();
    launch_tape_dump(docs, s.n, cap, d_ent.p, d_cnt.as<uint32_t>(), d_flg.as<uint32_t>(), c->stream);
    c->launches += 1;
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(entries, d_ent.p, (size_t)s.n * cap * 8, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(counts, d_cnt.p, (size_t)s.n * 4, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(flags, d_flg.p, (size_t)s.n * 4, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  d_ent.release();
  d_cnt.release();
  d_flg.release();
  if (e != cudaSuccess) return fail(c, "nbrm_tape_dump", e);
  return 0;
}

int nbrm_tensorboard_generate(nbrm_ctx* c, const nbrm_docs* tensorboards, const nbrm_docs* node_names,
                              uint32_t flags, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, tensorboards)) return r;
  if (node_names) {
    if (node_names->n != tensorboards->n) return fail_msg(c, "node_names must hold one document per Tensorboard");
    if (int r = nbrm_upload(c, 1, node_names)) return r;
  }
  if (int r = run_generator(c, GEN_ID_TENSORBOARD, {0, node_names ? 1 : -1, -1}, flags, nullptr, 0)) return r;
  return nbrm_download(c, out);
}

int nbrm_pvcviewer_reconcile(nbrm_ctx* c, const nbrm_docs* viewers, const nbrm_docs* found, const nbrm_docs* node_names,
                             nbrm_out* out) {
  if (!c || !out || !viewers) return -2;
  if (found && found->n != 3 * viewers->n) return fail_msg(c, "found must hold three documents per PVCViewer");
  if (node_names && node_names->n != viewers->n) return fail_msg(c, "node_names must hold one document per PVCViewer");
  if (int r = nbrm_upload(c, 0, viewers)) return r;
  nbrm_docs none;
  const uint64_t zero = 0;
  none.data = "";
  none.off = &zero;
  none.n = 0;
  if (int r = nbrm_upload(c, 1, found ? found : &none)) return r;  // the shared arena: empty = every Get NotFound
  if (node_names)
    if (int r = nbrm_upload(c, 2, node_names)) return r;
  if (int r = run_generator(c, GEN_ID_PVCVIEWER, {0, 1, node_names ? 2 : -1}, 0, nullptr, 0)) return r;
  return nbrm_download(c, out);
}

int nbrm_tensorboard_update(nbrm_ctx* c, const nbrm_docs* tensorboards, const nbrm_docs* desired,
                            const nbrm_docs* found, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, tensorboards)) return r;
  if (int r = nbrm_upload(c, 1, desired)) return r;
  if (int r = nbrm_upload(c, 2, found)) return r;
  if (int r = run_generator(c, GEN_ID_TB_UPDATE, {0, 1, 2}, 0, nullptr, 0)) return r;
  return nbrm_download(c, out);
}

int nbrm_pvcviewer_status(nbrm_ctx* c, const nbrm_docs* viewers, const nbrm_docs* deployments, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, viewers)) return r;
  if (int r = nbrm_upload(c, 1, deployments)) return r;
  if (int r = run_generator(c, GEN_ID_PVCVIEWER_STATUS, {0, 1, -1}, 0, nullptr, 0)) return r;
  return nbrm_download(c, out);
}

int nbrm_pvcviewer_webhook(nbrm_ctx* c, const nbrm_docs* viewers, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, viewers)) return r;
  if (int r = run_generator(c, GEN_ID_PVCVIEWER_DEFAULT, {0, -1, -1}, 0, nullptr, 0)) return r;
  return nbrm_download(c, out);
}

int nbrm_profile_generate_resident(nbrm_ctx* c) {
  if (!c) return -2;
  return run_generator(c, GEN_ID_PROFILE, {0, -1, -1}, 0, nullptr, 0);
}

int nbrm_profile_generate(nbrm_ctx* c, const nbrm_docs* profiles, nbrm_out* out) {
  if (!c || !out) return -2;
  if (int r = nbrm_upload(c, 0, profiles)) return r;
  if (int r = nbrm_profile_generate_resident(c)) return r;
  return nbrm_download(c, out);
}

}  // extern "C"
