// Stable LSD radix sort of (uint64 key, uint32 payload) pairs. See radix_sort.cuh.
#include "radix_sort.cuh"

namespace wh {

__global__ void sort_plan_reset_kernel(SortPlan* plan) {
  plan->bits_or = 0ull;
  plan->bits_and = ~0ull;
  plan->num_active = 0;
  plan->out = 0;
}

// Which bytes differ anywhere in the input? Those are the passes that must run.
__global__ void sort_plan_kernel(SortPlan* plan) {
  unsigned long long diff = plan->bits_or ^ plan->bits_and;
  int k = 0;
  for (int d = 0; d < 8; ++d)
    if ((diff >> (8 * d)) & 0xffull) plan->digit[k++] = d;
  plan->num_active = k;
  plan->out = k & 1;
}

// Per-block digit counts of pass p: counts[digit * nblocks + block].
__global__ void __launch_bounds__(kSortThreads)
sort_count_kernel(SortBufs b, int p, size_t n, unsigned nblocks) {
  if (p >= b.plan->num_active) return;
  __shared__ unsigned hist[256];
  const uint64_t* keys = b.keys[p & 1];
  const int shift = b.plan->digit[p] * 8;
  hist[threadIdx.x] = 0;
  __syncthreads();
  const size_t base = (size_t)blockIdx.x * kSortTile;
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll 4
  for (int r = 0; r < kSortItems; ++r) {
    size_t idx = base + (size_t)warp * (kSortItems * 32) + r * 32 + lane;
    bool valid = idx < n;
    unsigned d = valid ? (unsigned)((keys[idx] >> shift) & 0xff) : 0u;
    unsigned peers = match_any_bits<8>(d, valid);
    if (valid && (__ffs(peers) - 1) == (int)lane) atomicAdd(&hist[d], __popc(peers));
  }
  __syncthreads();
  b.counts[(size_t)threadIdx.x * nblocks + blockIdx.x] = hist[threadIdx.x];
}

// Exclusive scan over the digit-major count table (one block).
__global__ void __launch_bounds__(1024) sort_scan_kernel(SortBufs b, int p, size_t len) {
  if (p >= b.plan->num_active) return;
  __shared__ unsigned warp_tot[32];
  unsigned* c = b.counts;
  const size_t per = (len + 1023) / 1024;
  const size_t lo = min(len, (size_t)threadIdx.x * per), hi = min(len, lo + per);
  unsigned sum = 0;
  for (size_t i = lo; i < hi; ++i) sum += c[i];
  // block exclusive scan of the per-thread sums
  unsigned incl = sum;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    unsigned w = warp_tot[lane], wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    warp_tot[lane] = wi - w;
  }
  __syncthreads();
  unsigned run = warp_tot[warp] + incl - sum;
  for (size_t i = lo; i < hi; ++i) { unsigned t = c[i]; c[i] = run; run += t; }
}

// Stable scatter of pass p. Warp w owns a contiguous run of the tile and walks it in order, so
// "rank among equal digits" = (earlier warps' count) + (earlier rounds of this warp) + (lower lanes).
__global__ void __launch_bounds__(kSortThreads)
sort_scatter_kernel(SortBufs b, int p, size_t n, unsigned nblocks) {
  if (p >= b.plan->num_active) return;
  __shared__ unsigned cnt[kSortWarps][256];
  const uint64_t* __restrict__ keys = b.keys[p & 1];
  const uint32_t* __restrict__ vals = b.vals[p & 1];
  uint64_t* __restrict__ okeys = b.keys[(p & 1) ^ 1];
  uint32_t* __restrict__ ovals = b.vals[(p & 1) ^ 1];
  const int shift = b.plan->digit[p] * 8;
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int w = 0; w < kSortWarps; ++w) cnt[w][threadIdx.x] = 0;
  __syncthreads();

  const size_t base = (size_t)blockIdx.x * kSortTile + (size_t)warp * (kSortItems * 32) + lane;
  uint64_t key[kSortItems];
  uint32_t val[kSortItems];
  unsigned rank[kSortItems];
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    size_t idx = base + r * 32;
    bool valid = idx < n;
    key[r] = valid ? keys[idx] : 0;
    val[r] = valid ? vals[idx] : 0;
  }
  const unsigned lt = lanemask_lt();
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    bool valid = (base + r * 32) < n;
    unsigned d = valid ? (unsigned)((key[r] >> shift) & 0xff) : 0u;
    unsigned peers = match_any_bits<8>(d, valid);
    unsigned old = valid ? cnt[warp][d] : 0;
    __syncwarp();
    if (valid && (__ffs(peers) - 1) == (int)lane) cnt[warp][d] = old + __popc(peers);
    __syncwarp();
    rank[r] = old + __popc(peers & lt);
  }
  __syncthreads();
  {  // thread d turns per-warp counts of digit d into global start offsets
    unsigned run = b.counts[(size_t)threadIdx.x * nblocks + blockIdx.x];
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) { unsigned t = cnt[w][threadIdx.x]; cnt[w][threadIdx.x] = run; run += t; }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    if ((base + r * 32) < n) {
      unsigned d = (unsigned)((key[r] >> shift) & 0xff);
      unsigned pos = cnt[warp][d] + rank[r];
      okeys[pos] = key[r];
      ovals[pos] = val[r];
    }
  }
}

size_t sort_counts_bytes(size_t n) { return 256 * div_up(n ? n : 1, kSortTile) * sizeof(uint32_t); }

int sort_plan_reset(whContext* ctx, SortPlan* plan) {
  WH_LAUNCH(sort_plan_reset_kernel, 1, 1, 0, ctx->stream, plan);
  return WH_OK;
}

int radix_sort_pairs(whContext* ctx, const SortBufs& bufs, size_t n) {
  WH_LAUNCH(sort_plan_kernel, 1, 1, 0, ctx->stream, bufs.plan);
  if (n == 0) return WH_OK;
  WH_REQUIRE(n < (size_t)UINT32_MAX, "radix sort: %zu pairs do not fit 32-bit positions", n);
  const unsigned nblocks = (unsigned)div_up(n, kSortTile);
  for (int p = 0; p < 8; ++p) {
    WH_LAUNCH(sort_count_kernel, nblocks, kSortThreads, 0, ctx->stream, bufs, p, n, nblocks);
    WH_LAUNCH(sort_scan_kernel, 1, 1024, 0, ctx->stream, bufs, p, (size_t)256 * nblocks);
    WH_LAUNCH(sort_scatter_kernel, nblocks, kSortThreads, 0, ctx->stream, bufs, p, n, nblocks);
  }
  return WH_OK;
}

}  // namespace wh
