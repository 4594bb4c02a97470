// Shared internals of libwormhole_b200.so: error plumbing, the per-GPU context with its
// scratch arena, and small device helpers. Not part of the public ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <string>
#include <vector>

#include "../../include/wormhole_b200.h"

namespace wh {

void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;
// WH_TIMELINE=<file>: every profiler mark of every lane with its time since the process's first mark (a debugging aid:
// where a pipelined step spends its time; the marks cost an event record each)
extern FILE* g_timeline;

#define WH_CUDA(expr)                                                                      \
  do {                                                                                     \
    cudaError_t e__ = (expr);                                                              \
    if (e__ != cudaSuccess) {                                                              \
      wh::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
      return WH_ERR_CUDA;                                                                  \
    }                                                                                      \
  } while (0)

#define WH_TRY(expr)               \
  do {                             \
    int r__ = (expr);              \
    if (r__ != WH_OK) return r__;  \
  } while (0)

#define WH_REQUIRE(cond, ...)        \
  do {                               \
    if (!(cond)) {                   \
      wh::set_error(__VA_ARGS__);    \
      return WH_ERR_INVALID;         \
    }                                \
  } while (0)

// Launch + count + check. Grid sizes of zero are skipped (empty inputs).
#define WH_LAUNCH(kernel, grid, block, smem, stream, ...)                      \
  do {                                                                         \
    if ((grid) > 0) {                                                          \
      kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);              \
      wh::g_launches.fetch_add(1, std::memory_order_relaxed);                  \
      WH_CUDA(cudaGetLastError());                                             \
    }                                                                          \
  } while (0)

static inline size_t div_up(size_t a, size_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t a, size_t b) { return div_up(a, b) * b; }

// A grow-only device buffer: reused across calls, reallocated only when too small.
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes);
  void release();
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct PinnedBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes);
  void release();
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

}  // namespace wh

namespace wh {
// Opt-in per-kernel timing with CUDA events on the launching stream (bench.py's roofline leg).
struct Profiler {
  bool on = false;
  std::vector<cudaEvent_t> ev;        // ev[i] recorded right after segment i-1's kernels
  std::vector<const char*> name;      // name[i]: what ran between ev[i] and ev[i+1]
  size_t used = 0;
  struct Acc { std::string name; double ms = 0; unsigned long long count = 0; };
  std::vector<Acc> acc;
};
}  // namespace wh

struct whContext {
  int device = 0;
  wh::Profiler prof;
  int num_sms = 148;
  size_t l2_bytes = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = true;
  bool avoid_coop = false;   // this context's kernels run beside other cooperative launches (sharded jobs, several lanes)
  // scratch shared by the stateless entry points (grow-only)
  wh::DevBuf directory;      // direct-address localizer: {bits, prefix} per 32 keys
  size_t directory_words = 0;  // words currently known to be all-zero
  wh::DevBuf scan_tmp;       // block sums of the directory scan
  wh::DevBuf scan_state;     // tile states of the one-launch scan (scan.cuh); zeroed when (re)allocated
  unsigned scan_epoch = 0;   // call counter of scan_state
  cudaEvent_t side_ev = nullptr;   // orders work handed to another context's stream after this one's (created on first use)
  wh::DevBuf sort_tmp[6];    // radix sort ping-pong + histograms
  wh::DevBuf auc;            // histogram block of the counting AUC (loss.cu); zeroed when allocated
  wh::DevBuf acc;            // int64 fixed-point accumulators of the transposed products
  wh::DevBuf small;          // device scalars (counts, plans, metric partials)
  wh::DevBuf stage[8];       // staging of host arguments
  wh::PinnedBuf pinned;      // host landing zone for small read-backs
};

namespace wh {

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// Marks the end of a named segment of device work (no-op unless profiling is enabled).
int prof_begin(whContext* ctx);
int prof_mark(whContext* ctx, const char* name);
int prof_collect(whContext* ctx);  // after a stream sync: folds event pairs into the accumulators
#define WH_PROF(ctx, name) do { if ((ctx)->prof.on) WH_TRY(wh::prof_mark((ctx), (name))); } while (0)

// Brings a host array to the device (stage buffer) or passes a device pointer through.
template <typename T>
int to_device(whContext* ctx, DevBuf& buf, const T* src, size_t n, int mem, const T** out) {
  if (mem == WH_DEVICE || src == nullptr) { *out = src; return WH_OK; }
  WH_TRY(buf.reserve((n ? n : 1) * sizeof(T)));
  if (n) WH_CUDA(cudaMemcpyAsync(buf.p, src, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  *out = buf.as<T>();
  return WH_OK;
}

// Output side of the same convention: compute into a stage buffer for host callers, copy back.
template <typename T>
int out_buffer(whContext* ctx, DevBuf& buf, T* user, size_t n, int mem, T** dev) {
  if (mem == WH_DEVICE) { *dev = user; return WH_OK; }
  WH_TRY(buf.reserve((n ? n : 1) * sizeof(T)));
  *dev = buf.as<T>();
  return WH_OK;
}
template <typename T>
int out_commit(whContext* ctx, T* user, const T* dev, size_t n, int mem) {
  if (mem == WH_HOST && n && user) WH_CUDA(cudaMemcpyAsync(user, dev, n * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  return WH_OK;
}

// ---- device helpers ---------------------------------------------------------------------
#ifdef __CUDACC__

// learn/base/localizer.h:16-26 ReverseBytes (nibble reversal)
__host__ __device__ __forceinline__ uint64_t reverse_nibbles(uint64_t x) {
  x = x << 32 | x >> 32;
  x = (x & 0x0000FFFF0000FFFFULL) << 16 | (x & 0xFFFF0000FFFF0000ULL) >> 16;
  x = (x & 0x00FF00FF00FF00FFULL) << 8 | (x & 0xFF00FF00FF00FF00ULL) >> 8;
  x = (x & 0x0F0F0F0F0F0F0F0FULL) << 4 | (x & 0xF0F0F0F0F0F0F0F0ULL) >> 4;
  return x;
}

// learn/base/localizer.h:106-128: the key transform for I = uint64.
// mode 0: id % max_key (general), 1: id & (max_key-1) (power of two), 2: nibble reversal.
struct KeyXform {
  uint64_t max_key;
  int mode;
  __host__ __device__ __forceinline__ uint64_t operator()(uint64_t id) const {
    if (mode == 1) return id & (max_key - 1);
    if (mode == 0) return id % max_key;
    return reverse_nibbles(id);
  }
};
static inline KeyXform make_xform(uint64_t max_key) {
  KeyXform x;
  x.max_key = max_key;
  if (max_key == UINT64_MAX) x.mode = 2;
  else x.mode = ((max_key & (max_key - 1)) == 0 && max_key != 0) ? 1 : 0;
  return x;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// Lanes holding the same BITS-bit value, by BITS ballots. The hardware match.any serialises over
// the distinct values in the warp (measured here: ~1 us per call on 32 distinct digits), which
// made it the cost of every radix pass; ballots are a fixed ~3 instructions per bit.
template <int BITS>
__device__ __forceinline__ unsigned match_any_bits(unsigned v, bool valid) {
  unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
  for (int b = 0; b < BITS; ++b) {
    const unsigned bit = (v >> b) & 1u;
    const unsigned m = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? m : ~m;
  }
  return peers;  // meaningful for valid lanes only
}

// Lanes of the warp that hold the same 32-bit key as the caller (always includes the caller).
// Exact and cheap: group by a 5-bit hash with 5 ballots, then compare with the hash group's
// lowest lane; a lane that differs from that leader acts alone (it may miss a partner, never
// joins a wrong one).
__device__ __forceinline__ unsigned group_equal_keys(unsigned key, bool valid) {
  const unsigned lane = threadIdx.x & 31;
  const unsigned h = (key * 0x9E3779B1u) >> 27;
  const unsigned hp = match_any_bits<5>(h, valid);
  const int leader = valid ? (__ffs(hp) - 1) : (int)lane;
  const unsigned lk = __shfl_sync(0xffffffffu, key, leader);
  const bool same = valid && lk == key;
  const unsigned sm = __ballot_sync(0xffffffffu, same);
  return same ? (hp & sm) : (1u << lane);
}

// Sum of a 64-bit integer over the lanes in `group` (all of which must call with the same mask):
// four hardware warp reductions on 16-bit pieces, recombined modulo 2^64.
__device__ __forceinline__ long long group_sum_i64(unsigned group, long long v) {
  const unsigned long long u = (unsigned long long)v;
  const unsigned p0 = __reduce_add_sync(group, (unsigned)(u & 0xffffu));
  const unsigned p1 = __reduce_add_sync(group, (unsigned)((u >> 16) & 0xffffu));
  const unsigned p2 = __reduce_add_sync(group, (unsigned)((u >> 32) & 0xffffu));
  const unsigned p3 = __reduce_add_sync(group, (unsigned)(u >> 48));
  return (long long)((unsigned long long)p0 + ((unsigned long long)p1 << 16) + ((unsigned long long)p2 << 32) +
                     ((unsigned long long)p3 << 48));
}

// ---- TMA 1-D bulk copies (cp.async.bulk, SASS UBLKCP) with mbarrier completion ----------------
// A single thread hands the copy engine a whole tile; the bytes land in shared memory while the
// block works on earlier tiles. Memory-level parallelism then no longer depends on how many
// loads each thread can keep in flight, which is what limited the id-streaming kernels.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// global -> shared; dst, src 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "WH_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra WH_DONE;\n"
      "bra WH_WAIT;\n"
      "WH_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// Streaming loads that do not pollute L1 (inputs read once).
__device__ __forceinline__ uint64_t ld_stream_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}
// ... and that are the first to leave L2 as well (a stream of ids must not push the table entries out of L2)
__device__ __forceinline__ uint64_t ld_stream_evict_first_u64(const uint64_t* p) {
  uint64_t v, pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ uint64_t ld_stream_evict_first_u64(const uint64_t* p, uint64_t pol) {
  uint64_t v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void st_stream_evict_first_u32(uint32_t* p, uint32_t v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(p), "r"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ uint32_t ld_stream_evict_first_u32(const uint32_t* p, uint64_t pol) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ uint32_t ld_stream_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

#endif  // __CUDACC__
}  // namespace wh
