// nbrm_kernels.cu — the kernels every generator shares: exclusive scan of the
// planned object lengths and the gather-copy that writes the packed output.
#include <stdio.h>

#include <mutex>

#include "nbrm_dev.h"

namespace nbrm {

// ---------------------------------------------------------------- scan
// lens[n] (u32) -> off[n+1] (u64), three small kernels: per-tile sums, a scan
// of the tile sums by one block, per-tile scan plus the tile's prefix.

static const int SCAN_THREADS = 256;
static const int SCAN_ITEMS = 4;
static const int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;  // 1024 lengths per block

__device__ __forceinline__ uint64_t warp_incl_scan(uint64_t v, int lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint64_t t = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= d) v += t;
  }
  return v;
}

// Inclusive scan across a 256-thread block; returns this thread's inclusive
// value and the block total through `total`.
__device__ __forceinline__ uint64_t block_incl_scan(uint64_t v, uint64_t* total) {
  __shared__ uint64_t wsum[SCAN_THREADS / 32];
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint64_t inc = warp_incl_scan(v, lane);
  if (lane == 31) wsum[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint64_t s = lane < SCAN_THREADS / 32 ? wsum[lane] : 0;
    s = warp_incl_scan(s, lane);
    if (lane < SCAN_THREADS / 32) wsum[lane] = s;
  }
  __syncthreads();
  uint64_t prefix = w ? wsum[w - 1] : 0;
  *total = wsum[SCAN_THREADS / 32 - 1];
  __syncthreads();
  return inc + prefix;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_sums(const uint32_t* __restrict__ lens,
                                                               uint32_t n,
                                                               uint64_t* __restrict__ tile_sums) {
  uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  uint64_t s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k)
    if (base + k < n) s += lens[base + k];
  uint64_t total;
  block_incl_scan(s, &total);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_prefix(uint64_t* __restrict__ tile_sums,
                                                                 uint32_t n_tiles) {
  __shared__ uint64_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (uint32_t start = 0; start < n_tiles; start += SCAN_THREADS) {
    uint32_t i = start + threadIdx.x;
    uint64_t v = i < n_tiles ? tile_sums[i] : 0;
    uint64_t total;
    uint64_t inc = block_incl_scan(v, &total);
    uint64_t carry = carry_s;
    if (i < n_tiles) tile_sums[i] = carry + inc - v;  // exclusive
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + total;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_write_offsets(
    const uint32_t* __restrict__ lens, uint32_t n, const uint64_t* __restrict__ tile_prefix,
    uint64_t* __restrict__ off) {
  uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  uint64_t s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    v[k] = base + k < n ? lens[base + k] : 0;
    s += v[k];
  }
  uint64_t total;
  uint64_t inc = block_incl_scan(s, &total);
  uint64_t run = tile_prefix[blockIdx.x] + inc - s;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    if (base + k < n) off[base + k] = run;
    run += v[k];
    if (base + k + 1 == n) off[n] = run;
  }
}

void launch_scan(const uint32_t* lens, uint32_t n, uint64_t* tile_tmp, uint64_t* off,
                 cudaStream_t st) {
  if (n == 0) {
    cudaMemsetAsync(off, 0, sizeof(uint64_t), st);
    return;
  }
  uint32_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  scan_tile_sums<<<tiles, SCAN_THREADS, 0, st>>>(lens, n, tile_tmp);
  scan_tile_prefix<<<1, SCAN_THREADS, 0, st>>>(tile_tmp, tiles);
  scan_write_offsets<<<tiles, SCAN_THREADS, 0, st>>>(lens, n, tile_tmp, off);
}

// One 8-byte value to pinned host memory, stored by the SM: a cudaMemcpy of it would
// queue in the device-to-host copy engine behind a download of hundreds of megabytes
// (nbrm_download_begin) and the host, which needs the value to size the output, with it.
__global__ void store_u64_kernel(const uint64_t* src, volatile uint64_t* host_dst) {
  *host_dst = *src;
  __threadfence_system();
}

void launch_store_u64(const uint64_t* src, uint64_t* host_dst, cudaStream_t st) {
  store_u64_kernel<<<1, 1, 0, st>>>(src, host_dst);
}

// Status / result words into buffers a download lane owns (same reason: no copy engine).
__global__ void copy_u32_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, uint32_t n) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    dst[i] = src[i];
}

void launch_copy_u32(const uint32_t* src, uint32_t* dst, uint32_t n, cudaStream_t st) {
  if (!n) return;
  const uint32_t blocks = (n + 255) / 256 < 592 ? (n + 255) / 256 : 592;
  copy_u32_kernel<<<blocks, 256, 0, st>>>(src, dst, n);
}

// nbrm_nbc_reconcile: a Notebook the generators rejected (the reference would have panicked
// on Containers[0] / Ports[0]) has no desired objects, so the compare stage sees "nothing
// desired, nothing to do" for its three pairs.  Its generator status goes onto those pairs.
__global__ void merge_status_kernel(const uint32_t* __restrict__ gen, uint32_t* __restrict__ pair,
                                    uint32_t n_pairs, uint32_t per) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pairs; i += gridDim.x * blockDim.x) {
    const uint32_t g = gen[i / per];
    if (g) pair[i] = g;
  }
}

void launch_merge_status(const uint32_t* gen, uint32_t* pair, uint32_t n_pairs, uint32_t per, cudaStream_t st) {
  if (!n_pairs) return;
  const uint32_t blocks = (n_pairs + 255) / 256 < 592 ? (n_pairs + 255) / 256 : 592;
  merge_status_kernel<<<blocks, 256, 0, st>>>(gen, pair, n_pairs, per);
}

size_t scan_tmp_bytes(uint32_t n) {
  return sizeof(uint64_t) * ((size_t)(n + SCAN_TILE - 1) / SCAN_TILE + 1);
}

// ---------------------------------------------------------------- emit
// Gather-copy of every planned object into the packed output.
//
// Objects here are a few hundred bytes assembled from segments of a few dozen
// bytes (template text alternating with names), so a warp spread over one
// object idles most of its lanes on per-segment bookkeeping.  emit_objects
// therefore gives each object to ONE thread, and makes the 32 lanes of a warp
// work on the same object kind of 32 consecutive input documents (objects are
// stored document-major, nobj per document), so that the lanes run the same
// segment structure in lockstep.  A thread reads each source through aligned
// 16-byte blocks and writes its object through aligned 8-byte words; only the
// bytes before the first and after the last aligned word are stored singly,
// because those words are shared with the neighbouring objects.
//
// Objects longer than BIG_OBJECT bytes (a Notebook with a huge annotation) are
// left to emit_big_objects, which spreads a warp over each of them.

static const uint32_t BIG_OBJECT = 16384;

struct Writer {
  char* o;
  unsigned long long acc;
  uint32_t fill;  // bytes in acc
  uint32_t head;  // single bytes still to store before o is 8-byte aligned
  __device__ __forceinline__ void init(char* p) {
    o = p;
    acc = 0;
    fill = 0;
    head = (uint32_t)((8u - ((uint32_t)(unsigned long long)p & 7u)) & 7u);
  }
  // append the nb (1..4) low bytes of v
  __device__ __forceinline__ void put(uint32_t v, uint32_t nb) {
    while (head && nb) {
      *o++ = (char)v;
      v >>= 8;
      --nb;
      --head;
    }
    if (!nb) return;
    if (nb < 4) v &= (1u << (8u * nb)) - 1u;
    acc |= (unsigned long long)v << (8u * fill);
    fill += nb;
    if (fill >= 8) {
      *reinterpret_cast<unsigned long long*>(o) = acc;
      o += 8;
      fill -= 8;
      acc = fill ? ((unsigned long long)v >> (8u * (nb - fill))) : 0ull;
    }
  }
  // append 8 bytes at once (only once the head bytes are out: o is 8-byte aligned then)
  __device__ __forceinline__ void put8(unsigned long long v) {
    const uint32_t sh = 8u * fill;
    *reinterpret_cast<unsigned long long*>(o) = acc | (v << sh);
    o += 8;
    acc = fill ? (v >> (64u - sh)) : 0ull;
  }
  __device__ __forceinline__ void finish() {
    for (uint32_t k = 0; k < fill; ++k) o[k] = (char)(acc >> (8u * k));
  }
};

__device__ __forceinline__ void copy_segment(const char* src, uint32_t len, Writer& W) {
  if (!len) return;
  unsigned long long a = (unsigned long long)src;
  const unsigned long long last = (a + len - 1) & ~15ull;  // the aligned block that holds the last byte
  uint4 w = *reinterpret_cast<const uint4*>(a & ~15ull);
  while (len) {
    // the next block is on its way while this one is written out (a thread's loads are
    // otherwise a chain: descriptor -> source block -> source block ...)
    const unsigned long long nb = (a & ~15ull) + 16;
    uint4 wn = w;
    if (nb <= last) wn = *reinterpret_cast<const uint4*>(nb);
    const uint32_t k = (uint32_t)a & 15u;
    const uint32_t avail = min(16u - k, len);
    if (k == 0 && avail == 16 && W.head == 0) {
      // a whole source block (the inside of a long segment, e.g. a copied PodSpec): two 8-byte
      // appends instead of four 4-byte ones
      W.put8((unsigned long long)w.x | ((unsigned long long)w.y << 32));
      W.put8((unsigned long long)w.z | ((unsigned long long)w.w << 32));
      a += 16;
      len -= 16;
      w = wn;
      continue;
    }
    uint32_t done = 0;
    while (done < avail) {
      const uint32_t kk = k + done;
      const uint32_t word = (kk & 8u) ? ((kk & 4u) ? w.w : w.z) : ((kk & 4u) ? w.y : w.x);
      const uint32_t sh = kk & 3u;
      const uint32_t nb4 = min(4u - sh, avail - done);
      W.put(word >> (8u * sh), nb4);
      done += nb4;
    }
    a += avail;
    len -= avail;
    w = wn;
  }
}

__global__ void __launch_bounds__(EMIT_THREADS) emit_objects(const Seg* __restrict__ segs,
                                                            const uint32_t* __restrict__ nseg,
                                                            const uint32_t* __restrict__ lens,
                                                            const uint64_t* __restrict__ off,
                                                            uint32_t n_out, uint32_t cap,
                                                            uint32_t nobj, SrcBases bases,
                                                            char* __restrict__ out) {
  // thread -> object: lanes of a warp take the same kind k of 32 consecutive documents
  const uint32_t t = blockIdx.x * EMIT_THREADS + threadIdx.x;
  const uint32_t per_group = 32u * nobj;
  const uint32_t group = t / per_group, r = t - group * per_group;
  const uint32_t k = r >> 5, lane = r & 31u;
  const uint32_t obj = (group * 32u + lane) * nobj + k;
  if (obj >= n_out) return;
  const uint32_t ns = nseg[obj];
  if (!ns || lens[obj] > BIG_OBJECT) return;
  const Seg* s = segs + (size_t)obj * cap;
  Writer W;
  W.init(out + off[obj]);
  Seg next = s[0];
  for (uint32_t j = 0; j < ns; ++j) {
    const Seg sg = next;
    if (j + 1 < ns) next = s[j + 1];  // the next descriptor loads while this segment is copied
    copy_segment(bases.p[sg.len_src >> 28] + sg.off, sg.len_src & SEG_LEN_MASK, W);
  }
  W.finish();
}

// One warp per object, lanes stride over each segment's bytes; only for the
// rare objects emit_objects skips.
__global__ void __launch_bounds__(EMIT_THREADS) emit_big_objects(const Seg* __restrict__ segs,
                                                                const uint32_t* __restrict__ nseg,
                                                                const uint32_t* __restrict__ lens,
                                                                const uint64_t* __restrict__ off,
                                                                uint32_t n_out, uint32_t cap,
                                                                SrcBases bases,
                                                                char* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const uint32_t warps_per_block = EMIT_THREADS / 32;
  uint32_t w = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
  const uint32_t stride = gridDim.x * warps_per_block;
  for (; w < n_out; w += stride) {
    if (lens[w] <= BIG_OBJECT) continue;
    const uint32_t ns = nseg[w];
    const Seg* s = segs + (size_t)w * cap;
    char* o = out + off[w];
    for (uint32_t k = 0; k < ns; ++k) {
      const Seg sg = s[k];
      const uint32_t len = sg.len_src & SEG_LEN_MASK;
      const char* src = bases.p[sg.len_src >> 28] + sg.off;
      for (uint32_t j = lane; j < len; j += 32) o[j] = src[j];
      o += len;
    }
  }
}

void launch_emit(const Seg* segs, const uint32_t* nseg, const uint32_t* lens, const uint64_t* off,
                 uint32_t n_out, uint32_t cap, uint32_t nobj, const SrcBases& bases, char* out,
                 int sm_count, cudaStream_t st) {
  if (!n_out) return;
  // round the thread count up to whole (32 * nobj) groups
  const uint32_t per_group = 32u * nobj;
  const uint64_t threads = (uint64_t)((n_out / nobj + 31u) / 32u) * per_group;
  const uint32_t blocks = (uint32_t)((threads + EMIT_THREADS - 1) / EMIT_THREADS);
  emit_objects<<<blocks, EMIT_THREADS, 0, st>>>(segs, nseg, lens, off, n_out, cap, nobj, bases, out);
  uint32_t wpb = EMIT_THREADS / 32;
  uint32_t big_blocks = (n_out + wpb - 1) / wpb;
  uint32_t max_blocks = (uint32_t)sm_count * 8;
  if (big_blocks > max_blocks) big_blocks = max_blocks;
  emit_big_objects<<<big_blocks, EMIT_THREADS, 0, st>>>(segs, nseg, lens, off, n_out, cap, bases, out);
}

}  // namespace nbrm

// ---------------------------------------------------------------- scan without allocation
// nbrm_scan_lengths_noalloc (include/nbrm.h): the same three scan kernels, with
// the per-tile sums kept in a fixed array that is part of the loaded module
// instead of a cudaMalloc'ed buffer.  4096 tiles x 8 B = 32 KiB per GPU, held
// from module load until the process exits; one tile covers 1024 lengths, so
// one call scans at most 4 194 304 lengths.
//
// In host code the name g_noalloc_tile_sums is only a handle for the symbol,
// not an address a kernel can use: the device address is looked up once per
// device with cudaGetSymbolAddress (below) and that pointer is what the
// kernels get.
//
// All calls on a device share the one array.  Every slot a call reads (tiles
// 0..t-1) is written by scan_tile_sums earlier in the same call, so nothing
// has to start at zero and nothing is cleared; slots past t keep what an
// earlier call left and are never read.  Two calls whose kernels can run at
// the same time (different streams, no dependency between them) would
// overwrite each other's tile sums: the caller must order them.
static const uint32_t NOALLOC_MAX_TILES = 4096;
static __device__ __align__(256) uint64_t g_noalloc_tile_sums[NOALLOC_MAX_TILES];

namespace {
const int NOALLOC_MAX_DEVICES = 64;
uint64_t* g_noalloc_dev_ptr[NOALLOC_MAX_DEVICES];  // device address of the array, per device
std::mutex g_noalloc_mu;

int noalloc_fail(char* err, size_t cap, int rc, const char* what, const char* detail) {
  if (err && cap)
    snprintf(err, cap, "nbrm_scan_lengths_noalloc: %s%s%s", what, detail ? ": " : "",
             detail ? detail : "");
  return rc;
}
}  // namespace

extern "C" int nbrm_scan_lengths_noalloc(const uint32_t* d_lens, uint64_t n, uint64_t* d_off,
                                         void* stream, char* err, size_t err_cap) {
  if (err && err_cap) err[0] = 0;
  if (!d_off || (n && !d_lens)) return noalloc_fail(err, err_cap, -2, "null pointer", nullptr);
  const uint64_t max_n = (uint64_t)NOALLOC_MAX_TILES * nbrm::SCAN_TILE;
  if (n > max_n) {
    char msg[128];
    snprintf(msg, sizeof msg, "n = %llu but the fixed scratch covers at most %llu lengths per call",
             (unsigned long long)n, (unsigned long long)max_n);
    return noalloc_fail(err, err_cap, -3, "too many lengths", msg);
  }
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return noalloc_fail(err, err_cap, -1, "cudaGetDevice", cudaGetErrorString(e));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES)
    return noalloc_fail(err, err_cap, -2, "device ordinal out of range", nullptr);
  uint64_t* tiles = nullptr;
  {
    // Set-up: the device address of the array, looked up with that device current (it is
    // current here: cudaGetDevice just reported it).  The first call on a device is the one
    // that may load the module; later calls repeat the lookup, which is then a table read
    // and allocates nothing, because an address remembered from before a cudaDeviceReset
    // would point into a context that no longer exists.
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    uint64_t* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_tile_sums);
    if (e != cudaSuccess || !p)
      return noalloc_fail(err, err_cap, -1, "cudaGetSymbolAddress", cudaGetErrorString(e));
    g_noalloc_dev_ptr[dev] = p;
    tiles = p;
  }
  nbrm::launch_scan(d_lens, (uint32_t)n, tiles, d_off, (cudaStream_t)stream);
  e = cudaGetLastError();
  if (e != cudaSuccess) return noalloc_fail(err, err_cap, -1, "kernel launch", cudaGetErrorString(e));
  return 0;
}
