// whBatch: a localized minibatch living in HBM.
#pragma once
#include "common.cuh"
#include "radix_sort.cuh"

struct whBatch {
  whContext* ctx = nullptr;
  size_t rows = 0;
  size_t nnz = 0;        // entries of the localized block (after any dictionary drop)
  size_t nuniq = 0;      // unique transformed keys of the input block
  bool has_value = false, has_label = false, has_cnt = false;
  wh::DevBuf uniq;       // uint64[nuniq]
  wh::DevBuf cnt;        // uint32[nuniq]
  wh::DevBuf offset;     // size_t[rows+1]
  wh::DevBuf index;      // uint32[nnz]
  wh::DevBuf value;      // float[nnz]
  wh::DevBuf label;      // float[rows]
  ~whBatch() { uniq.release(); cnt.release(); offset.release(); index.release(); value.release(); label.release(); }
};

namespace wh {

// Device-side inputs of one localize call.
struct LocInput {
  const uint64_t* index;   // nnz raw ids
  size_t nnz;
  KeyXform xf;
};

// Direct-addressed localizer (hashed key spaces): outputs uniq (ascending), local ids, optional
// counts; *d_nuniq receives U. All stream-ordered, no host sync.
int localize_direct(whContext* ctx, const LocInput& in, uint64_t* d_uniq, uint32_t* d_cnt,
                    uint32_t* d_index, unsigned* d_nuniq);
// Radix-sort localizer (any 64-bit keys).
int localize_sort(whContext* ctx, const LocInput& in, uint64_t* d_uniq, uint32_t* d_cnt,
                  uint32_t* d_index, unsigned* d_nuniq);
bool direct_supported(uint64_t max_key);
struct HotView;
// true if localize_direct_hot can take a PullAtEmit for this key space on this device (the cooperative scan fits)
bool pull_at_emit_supported(const whContext* ctx, uint64_t max_key);
// Handle::Pull folded into the scan's emit (a local, dense, unsharded shard): the thread that emits a unique key also
// reads its weight, so the separate pull launch disappears from the chain. `after` (may be NULL) is waited for on the
// stream before the scan: the event of the push that this pull has to see.
struct PullAtEmit {
  const void* table = nullptr;   // the shard's ShardEntry array; NULL = no pull here
  int algo = 0;
  uint64_t key_begin = 0, key_end = 0;
  float* w = nullptr;            // out: weight per unique key
  long long* counters = nullptr; // [1] counts keys outside the shard
  cudaEvent_t after = nullptr;
};
// Direct localizer with the worker's hot-key table (hot.cuh): d_codes[i] = kHotFlag | slot for table keys, the rank
// otherwise; also fills hot.rank for this minibatch.
// resolve = false stops after the scan: d_codes[i] of an entry that is not a table key is its transformed key (< 2^28),
// to be turned into prefix + popc by the reader of the codes from ctx->directory, which the caller must zero
// (max_key/32 uint2 words) before the next localize on this context.
int localize_direct_hot(whContext* ctx, const LocInput& in, const HotView& hot, uint64_t* d_uniq, uint32_t* d_codes,
                        unsigned* d_nuniq, bool resolve = true, const PullAtEmit* pull = nullptr);

}  // namespace wh
