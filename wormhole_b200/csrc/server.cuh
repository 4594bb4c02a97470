// Piece (4): the HBM-resident model shard and its update handles. Internal API.
#pragma once
#include "common.cuh"
#include "spmv.cuh"

namespace wh {
constexpr int kMaxShards = 8;
// Barrier classes: each has its own flag row and epoch counter, so that barriers of different classes may complete in
// any relative order (0: pull/turn barriers and the trailing one, 1: key lists posted, 2: pulls served, 3: gradients posted).
constexpr int kBarrierClasses = 4;
// The shards of one box as seen from one GPU: every rank's table / counters / barrier flags are mapped
// into this process (CUDA IPC over NVLink), so a kernel can address any shard directly.
// One shard's table: an array of 32-byte entries, one DRAM sector each. What bounds every kernel that touches the
// table is the NUMBER of random lines it touches, not their bytes (B200 serves ~20-35 G random lines/s; measured:
// with w in an array of its own and the optimizer state in another, the push went from 14.5 to 20 us and the
// shared-GPU step paid two lines per key), so everything one key ever needs sits in one sector: the weight, the
// optimizer state -- FTRL {z, sq_cum_grad}, AdaGrad {-, sq_cum_grad}, SGD nothing -- and the key's fixed-point
// gradient accumulator of the shared-GPU training step (direct.cu; zero between minibatches). There the forward
// gather brings the sector into L2, the accumulating RED and the push then find it there.
struct alignas(32) ShardEntry {
  float w;           // SGDEntry / AdaGradEntry / FTRLEntry::w (learn/linear/async_sgd.h:72,108,146)
  float z, n;        // FTRLEntry::z / sq_cum_grad (:147-148), AdaGradEntry::sq_cum_grad (:109)
  unsigned pad0;
  long long acc;     // X^T dual of the minibatch in flight, fixed point
  long long pad1;
};
// A worker's outbox (one per lane; CUDA IPC, read and written by the owners of the keys):
//   header 256 bytes: word 0 = U (unique keys), words 4 .. 4+world = slice cuts, byte 128 = OutboxScale
//   keys   u64[cap]   ascending unique transformed keys
//   grads  8 bytes per key: the fp32 gradient in the low half, or (OutboxScale::fixed) the int64 fixed-point sum of
//          X^T dual exactly as the step's kernel left it -- the owner converts while it applies, which takes a launch
//          out of the push -> pull cycle
//   inbox  f32[cap]   the pulled weights, written by the owners
struct OutboxScale { double inv; float mul; unsigned fixed; };
__host__ __device__ static inline size_t outbox_keys_at() { return 256; }
__host__ __device__ static inline size_t outbox_grads_at(size_t cap) { return 256 + cap * 8; }
__host__ __device__ static inline size_t outbox_inbox_at(size_t cap) { return 256 + cap * 16; }
static inline size_t outbox_bytes(size_t cap) { return 256 + cap * 20; }
constexpr size_t kOutboxScaleAt = 128;

struct ShardView {
  ShardEntry* e = nullptr;     // [entries]
};
#if defined(__CUDACC__)
__host__ __device__
#endif
static inline ShardView shard_view(void* base) { ShardView v; v.e = static_cast<ShardEntry*>(base); return v; }
static inline size_t shard_bytes(size_t entries) { return entries * sizeof(ShardEntry); }
struct ShardMap {
  int world = 1, rank = 0;
  uint64_t bounds[kMaxShards + 1] = {0};          // shard g owns keys [bounds[g], bounds[g+1])
  void* table[kMaxShards] = {nullptr};              // shard g's ShardEntry array
  long long* counters[kMaxShards] = {nullptr};
  unsigned long long* flags[kMaxShards] = {nullptr};   // flags[g][cls * kMaxShards + r]: last epoch of class cls rank r announced to rank g
};
}  // namespace wh

struct whServer {
  whContext* ctx = nullptr;
  wh::ShardMap map;                 // world == 1: just this shard
  wh::DevBuf flags;                 // this rank's barrier flags (kBarrierClasses x kMaxShards x u64)
  unsigned long long epoch[wh::kBarrierClasses] = {0, 0, 0, 0};   // barriers issued so far, per class (all ranks in lock step)
  cudaEvent_t barrier_ev[wh::kBarrierClasses] = {nullptr, nullptr, nullptr, nullptr};   // keeps a class's barriers in issue order across streams
  unsigned long long steps = 0;     // training minibatches pushed so far (SGD: t of every shard follows from it)
  unsigned long long active_sources = ~0ull;   // bit r: rank r sends a push message in the next step (wh_server_set_active_sources)
  void* peer_base[3 * wh::kMaxShards] = {nullptr};   // pointers opened with cudaIpcOpenMemHandle
  int algo = WH_ALGO_FTRL;
  float l1 = 1.f, l2 = 0.f;
  float alpha = 0.1f, beta = 1.f;   // ISGDHandle defaults, learn/linear/async_sgd.h:49
  uint64_t key_begin = 0, key_end = 0;
  int t = 1;                        // SGDHandle::t, learn/linear/async_sgd.h:100
  unsigned long long version = 0;   // counts everything that changed the shard (push messages, set_state, load)
  wh::DevBuf table;                 // ShardEntry[entries]
  wh::ShardView view() const { return wh::shard_view(table.p); }
  wh::DevBuf counters;              // [0] new_w (long long), [1] out-of-range keys seen
  wh::DevBuf hkeys;                 // hash shard: the key held by each slot (all ones = free)
  uint64_t hslots = 0;              // 0 = dense shard
  size_t entries() const { return hslots ? (size_t)hslots : (size_t)(key_end - key_begin); }
};

namespace wh {
// All take device pointers; d_n (may be NULL) overrides n with a count held on the device.
int server_pull_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, float* vals);
int server_push_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, const float* grads);
// Gradients still in the fixed-point accumulator of the transposed product: converts, clears the
// accumulator and applies the update in one pass (no fp32 gradient array in between).
// Multi-shard versions (world > 1): pull gathers every key from the GPU that owns it; push_turn applies this
// worker's gradients to ONE shard, (rank + turn) % world; barrier is a stream-ordered rendezvous of all ranks.
int server_pull_sharded_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, float* vals);
int server_push_turn_launch(whServer* s, cudaStream_t stream, int turn, const uint64_t* keys, size_t n, const unsigned* d_n,
                            long long* acc, const FixedScale* fs, float mul, unsigned copies, size_t stride);
int server_barrier_launch(whServer* s, cudaStream_t stream, int cls = 0);
// Owner-served pull: for every source rank, the weights of the slice of its posted key list that this shard owns are
// gathered locally and written into that rank's inbox (behind keys and gradients in its outbox), coalesced over NVLink.
int server_serve_all_launch(whServer* s, cudaStream_t stream, unsigned char* const* outboxes, size_t cap);
// owner-side application of one source rank's outbox slice; cuts of an outbox for all destinations
// all sources of one step in one cooperative launch (outboxes[src], src = 0..world-1, ascending source order)
int server_apply_all_launch(whServer* s, cudaStream_t stream, unsigned char* const* outboxes, size_t cap);
int server_apply_from_launch(whServer* s, cudaStream_t stream, const unsigned char* outbox, size_t cap);
int outbox_cuts_launch(whServer* s, cudaStream_t stream, unsigned char* outbox);
int outbox_clear_launch(whServer* s, cudaStream_t stream, unsigned char* outbox, size_t cap);
int server_push_acc_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, long long* acc,
                           const FixedScale* fs, float mul, unsigned copies, size_t stride);
}  // namespace wh
