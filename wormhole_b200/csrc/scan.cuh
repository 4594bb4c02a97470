// Multi-block exclusive scan building blocks (tile sums -> scan of sums -> apply), templated on
// a functor that yields the value of element i so that flags never have to be materialised.
#pragma once
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

}  // namespace wh
