// Hot keys of a worker's minibatch stream.
//
// Feature ids follow a power law: at the Criteo shape 4 096 of the 256 000 keys of a minibatch carry 78 % of its
// 3.9 M entries, and they are the same keys minibatch after minibatch. Every entry costs one random L2 access in
// each of mark, rank, forward gather and backward atomic, and it is the number of those accesses (not their bytes)
// that bounds the step. So the worker keeps a small table of the stream's most frequent keys and serves them from
// shared memory / L1 instead:
//   * the table is a 4-way bucketed hash of the transformed keys (2 048 buckets x 4 tags, 32 KB), built on the host from a
//     census of one of the worker's minibatches (the first, then every WH_HOT_REFRESH-th) and copied into shared
//     memory by the kernels;
//   * an entry whose key is in the table gets the CODE  kHotFlag | slot  instead of its rank as local index; the
//     forward product reads its weight from a 32 KB slot-indexed array, the backward product adds it up in a
//     shared-memory accumulator and flushes the block's sums once (spmv.cu, spmv_fused_hot_kernel);
//   * everything else (presence bits, ranks, the ascending key list, pull, push) is as before, so the results do
//     not depend on which keys are in the table -- it is a cache, not a model. Integer accumulation keeps the
//     gradients bit-identical with and without it.
#pragma once
#include "common.cuh"

namespace wh {

constexpr int kHotBuckets = 2048, kHotWays = 4, kHotSlots = kHotBuckets * kHotWays;
constexpr unsigned kHotFlag = 0x80000000u, kHotEmpty = 0xffffffffu, kHotNone = 0xffffffffu;

// Device-side view of one lane's hot-key state (all arrays have kHotSlots entries).
struct HotView {
  const uint4* tags;     // shared by the lanes: bucket b = tags[b] = 4 transformed keys (kHotEmpty = free way)
  unsigned* present;     // 1 if the slot's key occurs in the current minibatch (reset by hot_rank)
  unsigned* rank;        // rank of the slot's key in the current minibatch, kHotNone if absent
  float* w;              // pulled weight of the slot's key (evaluation passes; the fused training kernel gathers its own)
};

#ifdef __CUDACC__
__device__ __forceinline__ unsigned hot_bucket(unsigned k32) { return (k32 * 0x9E3779B1u) >> (32 - 11); }
// slot of key k32 in the (shared-memory) table, or -1
__device__ __forceinline__ int hot_probe(const uint4* s_tags, unsigned k32) {
  const unsigned b = hot_bucket(k32);
  const uint4 t = s_tags[b];
  // which way holds the key: a 4-bit match mask and its lowest set bit (a chain of selects compiled to 9 instructions
  // per way; this is 4 compares, 3 inserts and one find-first-set). No way matches: ffs(0) - 1 = -1.
  const unsigned m = (t.x == k32 ? 1u : 0u) | (t.y == k32 ? 2u : 0u) | (t.z == k32 ? 4u : 0u) | (t.w == k32 ? 8u : 0u);
  return m ? (int)(b * kHotWays) + (__ffs(m) - 1) : -1;
}
__device__ __forceinline__ void hot_load_tags(uint4* s_tags, const uint4* g_tags) {
  for (int i = threadIdx.x; i < kHotBuckets; i += blockDim.x) s_tags[i] = __ldg(g_tags + i);
}
#endif

// Host side: chooses the table from (key, count) pairs -- most frequent first, a key whose bucket is full is left
// out -- and writes the 32 KB image. Returns the number of keys placed.
int hot_build_image(const uint64_t* keys, const uint32_t* counts, size_t n, uint32_t* image /* kHotSlots */);

}  // namespace wh
