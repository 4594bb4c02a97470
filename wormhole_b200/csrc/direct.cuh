// The worker's training step when worker and server share a GPU (1 worker, 1 server: the configuration the
// benchmark line and the parity bar are quoted in). Internal API; see direct.cu.
#pragma once
#include "common.cuh"
#include "loss.cuh"
#include "server.cuh"
#include "spmv.cuh"

namespace wh {

// The key table of this path: 1 024 buckets x 4 ways (16 KB of tags in shared memory). Smaller than the table of the
// chain of launches (hot.cuh): every block also keeps a 64-bit accumulator and a weight per slot in shared memory.
constexpr int kDxBuckets = 1024, kDxWays = 4, kDxSlots = kDxBuckets * kDxWays;
constexpr unsigned kDxEmpty = 0xffffffffu;
constexpr int kDxMaxBlocks = 1280;  // blocks of the step's kernel (one per SM and wave): rows of the per-block partial sums

// The worker's key table (device arrays, shared by the worker's lanes).
struct DirectTable {
  DevBuf tags;        // uint4[kDxBuckets]: 4 transformed keys per bucket (kDxEmpty = free way)
  DevBuf hot_w;       // f32[kDxSlots]: the shard's weight of every slot's key (mirror kept by the push kernel)
  int placed = 0;
  double coverage = 0;   // share of the census minibatch's entries whose key is in the table
  bool ready = false;
  unsigned long long seen_version = 0;   // whServer::version the mirror is known to match
  void release() { tags.release(); hot_w.release(); }
};

// Per-lane scratch. The bitmap is all zero between minibatches; the partial sums are overwritten by every step.
struct DirectLane {
  // (keys outside the table accumulate in the shard itself: ShardEntry::acc, server.cuh)
  DevBuf bitmap;       // u32[max_key / 32]: keys outside the table that occur in the minibatch
  DevBuf part;         // int64[kDxMaxBlocks][kDxSlots]: each block's sums of X^T dual for the table keys
  DevBuf part_touched; // u32[kDxMaxBlocks][kDxSlots / 32]: ... and which table keys the block saw
  DevBuf codes;        // u32[nnz]: the minibatch's codes when they are made ahead of the step (direct_stage_launch)
  DevBuf touched;      // u32[kDxSlots]: table keys seen by direct_stage_launch; cleared by the push
  uint64_t keys = 0;
  unsigned blocks = 0; // rows of `part` the last step wrote
  void release() { bitmap.release(); part.release(); part_touched.release(); codes.release(); touched.release(); }
};

// Builds the table image from one minibatch's (key, count) pairs (host arrays) and uploads it; synchronous.
int direct_table_build(whContext* ctx, DirectTable* t, const uint64_t* keys, const uint32_t* counts, size_t n);
// hot_w[slot] = w[key of slot] for every slot (stream-ordered): after the table is built, and whenever the shard was
// changed by anything other than this worker's own push.
int direct_table_refresh_w(whContext* ctx, DirectTable* t, const ShardEntry* shard);
int direct_lane_prepare(whContext* ctx, DirectLane* l, uint64_t max_key);

struct DirectStep {
  int loss; bool train;
  size_t rows, nnz;
  const size_t* offset; const uint64_t* index; const float* value; const float* label;
  KeyXform xf;
  ShardEntry* shard;            // the shard's entries, addressed by key (key_begin == 0); accumulators are added into
  const uint32_t* codes;        // NULL, or the codes direct_stage_launch made for this minibatch
  float* xw;                    // out: X w per row
  double* part_objv; unsigned* part_correct;   // out: per-block evaluation partial sums (grid entries)
  unsigned long long* auc_hist; // the AUC histogram every row is added to
  long long* counters;          // the server's counters ([1]: errors reported by wh_worker_wait)
  FixedScale* fs;               // device copy of the fixed-point scale (valued features: written by fixed_prepare)
};
// Localize + Pull + Init + Evaluate + CalcGrad of AsgdWorker::ProcessMinibatch in one launch. Returns the grid size
// (= number of evaluation partials) in *nparts.
int direct_fused_launch(whContext* ctx, const DirectStep& st, const DirectTable& t, DirectLane& l, unsigned* nparts);
// The part of that launch that does not depend on the model -- key transform, table lookup, key bitmap -- as a launch
// of its own, for workers with several minibatches in flight: it runs beside the step of the minibatch before. The
// fused launch then takes st.codes = l.codes.
int direct_stage_launch(whContext* ctx, const DirectStep& st, const DirectTable& t, DirectLane& l);
// ZPush: one Handle::Push per key that occurred, gradients taken from the lane's accumulators (which it clears).
int direct_push_launch(whServer* s, cudaStream_t stream, const DirectTable& t, const DirectLane& l, const FixedScale* fs,
                       float mul, int num_sms);

}  // namespace wh
