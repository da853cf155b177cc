// nbrm_kernels.cu — the kernels every generator shares: exclusive scan of the
// planned object lengths and the gather-copy that writes the packed output.
#include <stdio.h>

#include <mutex>

#include "nbrm_dev.h"
#include "nbrm_emit.cuh"

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
// Gather-copy of every planned object into the packed output: one warp per object, one lane
// per 16-byte block of the output (nbrm_emit.cuh has the per-block work and the reasons).
//
// The segment list of an object is taken in windows of EMITW_WIN segments: the warp loads a
// window with coalesced 8-byte loads, scans the lengths with shuffles and leaves segments and
// prefix sums in shared memory.  The output is walked in rounds of 32 aligned 16-byte blocks;
// nbrm_emit.cuh says how a round is cut into pieces and dealt to the lanes.  Objects of any
// length and any number of segments take this path.

static const uint32_t EMITW_WARPS = EMIT_THREADS / 32;
static const uint32_t EMITW_WIN = 128;

__global__ void __launch_bounds__(EMIT_THREADS) emit_warp(const Seg* __restrict__ segs,
                                                         const uint32_t* __restrict__ nseg,
                                                         const uint64_t* __restrict__ off,
                                                         uint32_t n_out, uint32_t cap, uint32_t opw,
                                                         SrcBases bases, char* __restrict__ out) {
  __shared__ uint32_t sP[EMITW_WARPS][EMITW_WIN + 1];
  __shared__ Seg sS[EMITW_WARPS][EMITW_WIN];
  __shared__ const char* sB[SRC_COUNT];
  __shared__ __align__(16) uint32_t sscr[EMITW_WARPS][32 * 12];  // per lane: the source words of its piece (nbrm_emit.cuh)
  __shared__ uint4 sT[17];                         // sT[x]: all bytes below position x set
  if (threadIdx.x < 17u)
    sT[threadIdx.x] = make_uint4(emitw::below_word(threadIdx.x, 0), emitw::below_word(threadIdx.x, 1),
                                 emitw::below_word(threadIdx.x, 2), emitw::below_word(threadIdx.x, 3));
  if (threadIdx.x == 0) {
#pragma unroll
    for (uint32_t k = 0; k < SRC_COUNT; ++k) sB[k] = bases.p[k];
  }
  __syncthreads();
  const uint32_t FULL = 0xffffffffu;
  const uint32_t lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
  // a warp takes `opw` consecutive objects: their segment counts and offsets come with one
  // coalesced load each, and objects that have nothing to write (a compare that found no
  // difference) cost nothing more than that
  const uint32_t base = (blockIdx.x * EMITW_WARPS + wib) * opw;
  if (base >= n_out) return;
  uint32_t my_ns = 0;
  unsigned long long my_off = 0;
  if (lane < opw && base + lane < n_out) {
    my_ns = nseg[base + lane];
    my_off = off[base + lane];
  }
  uint32_t* P = sP[wib];
  Seg* S = sS[wib];
  uint32_t* scr = sscr[wib];
  uint32_t live = __ballot_sync(FULL, my_ns != 0u);
  while (live) {
    const uint32_t oi = (uint32_t)__ffs((int)live) - 1u;
    live &= live - 1u;
    const uint32_t ns = __shfl_sync(FULL, my_ns, oi);
    const Seg* os = segs + (size_t)(base + oi) * cap;
    char* dst = out + __shfl_sync(FULL, my_off, oi);
    for (uint32_t w0 = 0; w0 < ns; w0 += EMITW_WIN) {
      const uint32_t m = min(EMITW_WIN, ns - w0);
      uint32_t run = 0;
      for (uint32_t j0 = 0; j0 < m; j0 += 32u) {
        const uint32_t j = j0 + lane;
        Seg s;
        s.off = 0;
        s.len_src = 0;
        if (j < m) s = os[w0 + j];
        uint32_t v = s.len_src & SEG_LEN_MASK;
#pragma unroll
        for (uint32_t d = 1; d < 32u; d <<= 1) {
          const uint32_t t = __shfl_up_sync(FULL, v, d);
          if (lane >= d) v += t;
        }
        if (j < m) {
          S[j] = s;
          P[j + 1] = run + v;
        }
        run += __shfl_sync(FULL, v, 31);
      }
      if (lane == 0) P[0] = 0;
      __syncwarp();
      const uint32_t L = run;  // bytes of this window
      const uint32_t head = (uint32_t)((unsigned long long)dst & 15ull);
      char* blk0 = dst - head;
      const uint32_t nblocks = (head + L + 15u) >> 4;
      uint32_t first = 0;  // the segment the round's first byte lies in
      for (uint32_t c = 0; c < nblocks; c += 32u) {
        const uint32_t nbr = min(32u, nblocks - c);
        const uint32_t qs = c ? 16u * c - head : 0u;
        const uint32_t qe = min(L, 16u * (c + nbr) - head);
        // segments that start inside the round (the prefix sums are ascending)
        uint32_t cnt = 0;
        for (;;) {
          const uint32_t t = first + 1u + cnt + lane;
          const uint32_t bal = __ballot_sync(FULL, t < m && P[t] < qe);
          cnt += __popc(bal);
          if (bal != FULL) break;
        }
        const bool ragged = (c == 0u && head) || 16u * (c + nbr) - head > L;  // a block shared with a neighbour
        char* rdst = blk0 + 16ull * c;
        if (cnt == 0u && !ragged) {
          // the round lies inside segment `first`: source to output, one shift for the whole warp
          const Seg sg = S[first];
          if (lane < nbr) emitw::copy_pure_block(sB[sg.len_src >> 28] + sg.off + (qs - P[first]), rdst, lane);
        } else if (lane < nbr) {
          const uint32_t b = c + lane;
          const uint32_t q = b ? 16u * b - head : 0u;
          const uint32_t g = cnt ? emitw::find_seg(P, first, first + cnt, q) : first;
          emitw::emit_block(P, S, g, sB, q, min(L, 16u * (b + 1u) - head), b ? 0u : head, rdst + 16ull * lane,
                            scr + 12u * lane, reinterpret_cast<const uint32_t*>(sT));
        }
        first += cnt;
        if (first + 1u < m && P[first + 1u] == qe) ++first;
      }
      dst += L;
      __syncwarp();
    }
  }
}

void launch_emit(const Seg* segs, const uint32_t* nseg, const uint32_t* lens, const uint64_t* off,
                 uint32_t n_out, uint32_t cap, uint32_t nobj, const SrcBases& bases, char* out,
                 int sm_count, cudaStream_t st) {
  (void)lens;
  (void)nobj;
  (void)sm_count;
  if (!n_out) return;
  // objects per warp: one while the batch is small (latency), up to eight once there are more
  // warps than the GPU holds at a time
  uint32_t opw = n_out / 8192u;
  opw = opw < 1u ? 1u : opw > 8u ? 8u : opw;
  const uint32_t warps = (n_out + opw - 1) / opw;
  const uint32_t blocks = (warps + EMITW_WARPS - 1) / EMITW_WARPS;
  emit_warp<<<blocks, EMIT_THREADS, 0, st>>>(segs, nseg, off, n_out, cap, opw, bases, out);
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
