// Multi-block exclusive scan building blocks (tile sums -> scan of sums -> apply), templated on
// a functor that yields the value of element i so that flags never have to be materialised.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace wh {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

// Block-wide exclusive scan of one value per thread; returns the exclusive prefix and the total.
// T is unsigned or unsigned long long (the latter lets two 32-bit counters ride in one scan).
template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T* total) {
  __shared__ T warp_tot[32];
  __shared__ T block_tot;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
  T incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    T t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  __syncthreads();  // protect warp_tot reuse across calls
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    T w = lane < nwarps ? warp_tot[lane] : 0, wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      T t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    if (lane < nwarps) warp_tot[lane] = wi - w;
    if (lane == 31) block_tot = wi;
  }
  __syncthreads();
  *total = block_tot;
  return warp_tot[warp] + incl - v;
}

// Exclusive scan, in place, of `len` block sums by ONE block; writes the grand total to *total.
template <typename T>
__global__ void __launch_bounds__(1024) scan_sums_kernel(T* sums, size_t len, T* total) {
  const size_t per = (len + blockDim.x - 1) / blockDim.x;
  const size_t lo = min(len, (size_t)threadIdx.x * per), hi = min(len, lo + per);
  T s = 0;
  for (size_t i = lo; i < hi; ++i) s += sums[i];
  T tot;
  T run = block_exclusive_scan<T>(s, &tot);
  for (size_t i = lo; i < hi; ++i) { T t = sums[i]; sums[i] = run; run += t; }
  if (threadIdx.x == 0 && total) *total = tot;
}

// sums[block] = sum over the block's tile of f(i)
template <typename T, typename F>
__global__ void __launch_bounds__(kScanThreads) scan_tile_sums_kernel(F f, size_t n, T* sums) {
  const size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
  T s = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) if (base + k < n) s += f(base + k);
  T tot;
  block_exclusive_scan<T>(s, &tot);
  if (threadIdx.x == 0) sums[blockIdx.x] = tot;
}

// Calls g(i, value_i, exclusive_prefix_i) for every element, prefixes being global.
template <typename T, typename F, typename G>
__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(F f, G g, size_t n, const T* sums) {
  const size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
  T v[kScanItems], s = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) { v[k] = (base + k < n) ? f(base + k) : 0; s += v[k]; }
  T tot;
  T run = sums[blockIdx.x] + block_exclusive_scan<T>(s, &tot);
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    if (base + k < n) g(base + k, v[k], run);
    run += v[k];
  }
}

// Host driver: sums must hold div_up(n, kScanTile) values; *total receives the grand total.
template <typename T, typename F, typename G>
int scan_and_apply(whContext* ctx, F f, G g, size_t n, T* sums, T* total) {
  const unsigned nb = (unsigned)div_up(n ? n : 1, kScanTile);
  WH_LAUNCH((scan_tile_sums_kernel<T, F>), nb, kScanThreads, 0, ctx->stream, f, n, sums);
  WH_LAUNCH((scan_sums_kernel<T>), 1, 1024, 0, ctx->stream, sums, (size_t)nb, total);
  WH_LAUNCH((scan_apply_kernel<T, F, G>), nb, kScanThreads, 0, ctx->stream, f, g, n, sums);
  return WH_OK;
}

// ---- the same scan in ONE launch (decoupled look-back) ------------------------------------------------------
// Three launches of a few microseconds each are mostly launch latency when the input is a 4 MB directory. Here a
// block takes a tile by ticket (so that every tile's predecessors are already running), publishes its tile sum,
// looks back over its predecessors' published {aggregate | inclusive prefix} words until it meets a prefix, and
// applies. State words carry an epoch, so nothing has to be cleared between calls:
//   word = epoch << 42 | status << 40 | value (40 bits),  status 1 = tile aggregate, 2 = inclusive prefix.
// 32-bit scanned values; totals must stay below 2^40.
struct ScanState {
  unsigned long long* word;   // one per tile
  unsigned* ticket;           // self-resetting tile counter (atomicInc wraps at the tile count)
  unsigned epoch;             // 1 .. 2^22-1, advanced per call by the host
};
__device__ __forceinline__ unsigned long long scan_ld(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
template <int ITEMS, typename F, typename G>
__global__ void __launch_bounds__(kScanThreads) scan_onepass_kernel(F f, G g, size_t n, ScanState st, unsigned ntiles,
                                                                    unsigned* total) {
  __shared__ unsigned s_tile;
  __shared__ unsigned long long s_prefix;
  if (threadIdx.x == 0) s_tile = atomicInc(st.ticket, ntiles - 1);
  __syncthreads();
  const unsigned tile = s_tile;
  const size_t base = (size_t)tile * (kScanThreads * ITEMS) + (size_t)threadIdx.x * ITEMS;
  unsigned v[ITEMS], s = 0;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) { v[k] = (base + k < n) ? f(base + k) : 0u; s += v[k]; }
  unsigned tot;
  const unsigned excl = block_exclusive_scan<unsigned>(s, &tot);
  const unsigned long long tag = (unsigned long long)st.epoch << 42;
  if (threadIdx.x == 0) {
    const unsigned long long w = tag | ((tile == 0 ? 2ull : 1ull) << 40) | tot;
    atomicExch(st.word + tile, w);
    if (tile == 0) s_prefix = 0;
  }
  if (tile != 0 && threadIdx.x < 32) {
    unsigned long long sum = 0;
    int j = (int)tile - 1 - (int)threadIdx.x;      // this lane's predecessor in the current window
    // (fetching several windows per round trip was slower: words read early are still aggregates, read late
    //  they are often prefixes that end the walk)
    while (true) {
      unsigned long long w = 0;
      bool have = j < 0;                            // before tile 0: contributes nothing, counts as a prefix
      if (!have) {
        do { w = scan_ld(st.word + j); } while ((w >> 42) != st.epoch);
      }
      const bool is_prefix = have || ((w >> 40) & 3ull) == 2ull;
      const unsigned pm = __ballot_sync(0xffffffffu, is_prefix);
      const int first = pm ? __ffs(pm) - 1 : 32;    // nearest predecessor that already knows its prefix
      unsigned long long c = (!have && (int)threadIdx.x <= first) ? (w & ((1ull << 40) - 1)) : 0ull;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
      sum += c;
      if (pm) break;
      j -= 32;
    }
    if (threadIdx.x == 0) {
      s_prefix = sum;
      __threadfence();
      atomicExch(st.word + tile, tag | (2ull << 40) | (sum + tot));
    }
  }
  __syncthreads();
  unsigned run = (unsigned)s_prefix + excl;
  if (threadIdx.x == 0 && tile == ntiles - 1 && total) *total = (unsigned)s_prefix + tot;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    if (base + k < n) g(base + k, v[k], run);
    run += v[k];
  }
}

// Host driver. `state` must hold scan_onepass_bytes(n) bytes that were zero when first used and are only ever
// passed to this function; *epoch is the caller's call counter for that buffer.
constexpr int kOnepassItems = 4;   // elements per thread: smaller tiles than the 3-launch scan, for more blocks in flight
static inline size_t scan_onepass_bytes(size_t n) { return div_up(n ? n : 1, kScanThreads * kOnepassItems) * 8 + 64; }
template <typename F, typename G>
int scan_onepass(whContext* ctx, F f, G g, size_t n, void* state, unsigned* epoch, unsigned* total) {
  const unsigned nt = (unsigned)div_up(n ? n : 1, kScanThreads * kOnepassItems);
  if (++*epoch >= (1u << 22)) {   // epoch wrap: start over with a clean buffer
    WH_CUDA(cudaMemsetAsync(state, 0, scan_onepass_bytes(n), ctx->stream));
    *epoch = 1;
  }
  ScanState st;
  st.ticket = reinterpret_cast<unsigned*>(state);
  st.word = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(state) + 64);
  st.epoch = *epoch;
  WH_LAUNCH((scan_onepass_kernel<kOnepassItems, F, G>), nt, kScanThreads, 0, ctx->stream, f, g, n, st, nt, total);
  return WH_OK;
}

// ---- and in one COOPERATIVE launch, for inputs whose tiles are all resident at once ------------------------------
// The look-back above is a chain: all tiles start together, none finds a finished prefix nearby, and the late tiles
// walk far back, one L2 round trip per 32 predecessors (14 us for the 4 MB directory). With every tile resident a grid
// barrier replaces the chain: tile sums, barrier, each block adds up its predecessors' sums in one strided read,
// apply -- the elements stay in registers across the barrier.
// H: optional extra phases, each separated from the scan by a grid barrier: h.before() runs in every block before any
// element is read, h() after every g(...) has been applied.
struct NoEpilogue {
  __device__ void before() const {}
  __device__ void operator()() const {}
  static constexpr bool kPresent = false;
};
template <int ITEMS, typename F, typename G, typename H>
__global__ void __launch_bounds__(kScanThreads) scan_coop_kernel(F f, G g, H h, size_t n, unsigned* sums, unsigned* total) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  if (H::kPresent) {
    h.before();
    __threadfence();
    grid.sync();
  }
  const unsigned tile = blockIdx.x;
  const size_t base = (size_t)tile * (kScanThreads * ITEMS) + (size_t)threadIdx.x * ITEMS;
  unsigned v[ITEMS], s = 0;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) { v[k] = (base + k < n) ? f(base + k) : 0u; s += v[k]; }
  unsigned tot;
  const unsigned excl = block_exclusive_scan<unsigned>(s, &tot);
  if (threadIdx.x == 0) __stcg(sums + tile, tot);
  __threadfence();
  grid.sync();
  unsigned p = 0;
  for (unsigned j = threadIdx.x; j < tile; j += kScanThreads) p += __ldcg(sums + j);
  unsigned prefix;
  block_exclusive_scan<unsigned>(p, &prefix);
  if (threadIdx.x == 0 && tile == gridDim.x - 1 && total) *total = prefix + tot;
  unsigned run = prefix + excl;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    if (base + k < n) g(base + k, v[k], run);
    run += v[k];
  }
  if (H::kPresent) {
    __threadfence();
    grid.sync();
    h();
  }
}
// true if the cooperative variant can take an input of n elements on this device (all tiles resident)
static inline bool scan_coop_fits(const whContext* ctx, size_t n) {
  return div_up(n ? n : 1, kScanThreads * kOnepassItems) <= (size_t)ctx->num_sms * 4;
}
template <typename F, typename G, typename H = NoEpilogue>
int scan_coop(whContext* ctx, F f, G g, size_t n, unsigned* sums, unsigned* total, H h = H()) {
  const unsigned nt = (unsigned)div_up(n ? n : 1, kScanThreads * kOnepassItems);
  void* args[] = {&f, &g, &h, &n, &sums, &total};
  WH_CUDA(cudaLaunchCooperativeKernel((const void*)scan_coop_kernel<kOnepassItems, F, G, H>, dim3(nt), dim3(kScanThreads), args, 0,
                                      ctx->stream));
  wh::g_launches.fetch_add(1, std::memory_order_relaxed);
  return WH_OK;
}

}  // namespace wh
