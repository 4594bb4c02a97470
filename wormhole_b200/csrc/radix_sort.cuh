// Stable LSD radix sort of (uint64 key, uint32 payload) pairs, 8-bit digits, multi-block.
//
// Used by the sort-based localizer (learn/base/localizer.h:130-131 replaced: the reference's
// thread-recursive merge sort of 12-byte Pair structs, learn/base/parallel_sort.h:15-41) and by
// the large-input AUC. Passes whose digit is constant over the whole input are skipped; which
// ones is decided ON THE DEVICE from the OR / AND of all keys, so the host never synchronises:
// it always enqueues 8 x (count, scan, scatter) and the kernels of skipped passes return at once.
// Where the sorted data ends up (buffer 0 or 1) is therefore also device-side knowledge:
// consumers take both buffers plus the SortPlan and pick with sorted_keys()/sorted_vals().
#pragma once
#include "common.cuh"

namespace wh {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 16;
constexpr int kSortTile = kSortThreads * kSortItems;  // 4096 pairs per block
constexpr int kSortWarps = kSortThreads / 32;

struct SortPlan {
  unsigned long long bits_or;   // OR of all keys   (producer fills these two; init 0 / ~0)
  unsigned long long bits_and;  // AND of all keys
  int num_active;               // number of passes that do work
  int digit[8];                 // digit index (0 = least significant byte) of active pass p
  int out;                      // buffer (0/1) holding the sorted result
};

struct SortBufs {
  uint64_t* keys[2];
  uint32_t* vals[2];
  uint32_t* counts;  // 256 * nblocks
  SortPlan* plan;
};

__device__ __forceinline__ const uint64_t* sorted_keys(const SortBufs& b) { return b.keys[b.plan->out]; }
__device__ __forceinline__ const uint32_t* sorted_vals(const SortBufs& b) { return b.vals[b.plan->out]; }

size_t sort_counts_bytes(size_t n);
// keys[0]/vals[0] hold the input; plan->bits_or/bits_and must already be final on the stream.
int radix_sort_pairs(whContext* ctx, const SortBufs& bufs, size_t n);
// Helper: plan init (bits_or = 0, bits_and = ~0) as a stream-ordered memset-like kernel.
int sort_plan_reset(whContext* ctx, SortPlan* plan);

}  // namespace wh
