// Piece (1): the feature-id localizer.
//
// Replaces Localizer<uint64_t>::{CountUniqIndex, RemapIndex, Localize}
// (learn/base/localizer.h:42-51,96-221) and the ParallelSort under it
// (learn/base/parallel_sort.h:15-41). Two device strategies, bit-identical outputs:
//
//  direct  For hashed key spaces (ps::FLAGS_max_key <= 2^28). The transformed key indexes a
//          directory of {presence bits, rank prefix} per 32 keys. One pass marks presence, a
//          scan over the directory ranks the words and emits the ascending unique list, one
//          more pass turns every id into its rank. No sort: per minibatch the ids are read
//          twice and the directory (max_key/4 bytes, L2 resident for 2^24) is touched at random.
//  sort    For arbitrary 64-bit keys (the reference's default: nibble-reversed ids). Stable LSD
//          radix sort of (key, position), run-length heads -> ranks, scatter by position.
#include <stdlib.h>

#include <algorithm>
#include <numeric>
#include <vector>

#include "batch.cuh"
#include "hot.cuh"
#include "scan.cuh"
#include "server.cuh"

namespace wh {

constexpr uint64_t kDirectMaxKey = 1ull << 28;
bool direct_supported(uint64_t max_key) { return max_key != 0 && max_key <= kDirectMaxKey; }

// ------------------------------------------------------------------------------- direct --

// learn/base/localizer.h:107-115 (key transform) fused with "is this key present".
// Feature ids follow a power law: a few thousand keys make up most of the entries. Each block keeps
// a direct-mapped cache of keys it has already marked in shared memory (exact: the full key is
// compared), so repeats cost one shared-memory probe instead of an L2 load of the same hot word
// from every SM. Misses test the directory word first; only new keys pay for an atomic.
constexpr int kMarkCache = 8192;
constexpr int kLocIlp = 4;   // entries per thread per trip: the loads of a trip are issued together
__global__ void __launch_bounds__(256) loc_mark_kernel(const uint64_t* __restrict__ index, size_t n,
                                                       KeyXform xf, uint2* dir) {
  __shared__ unsigned cache[kMarkCache];
  for (int i = threadIdx.x; i < kMarkCache; i += blockDim.x) cache[i] = 0xffffffffu;
  __syncthreads();
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += stride * kLocIlp) {
    uint64_t k[kLocIlp];
    unsigned word[kLocIlp];
    bool probe[kLocIlp];
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) {
      const size_t i = i0 + u * stride;
      probe[u] = i < n;
      k[u] = probe[u] ? ld_stream_u64(index + i) : 0;
    }
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) {
      k[u] = xf(k[u]);
      const unsigned k32 = (unsigned)k[u];            // direct addressing: keys < 2^28
      const unsigned slot = (k32 * 0x9E3779B1u) >> (32 - 13);
      if (probe[u]) {
        if (cache[slot] == k32) probe[u] = false;     // this block already marked it
        else cache[slot] = k32;                       // racy by design: a lost update only costs a re-probe
      }
    }
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) word[u] = probe[u] ? __ldcg(&dir[k[u] >> 5].x) : 0xffffffffu;
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) {
      const unsigned bit = 1u << (k[u] & 31);
      if (probe[u] && !(word[u] & bit)) atomicOr(&dir[k[u] >> 5].x, bit);
    }
  }
}

struct DirPopc {
  const uint2* dir;
  __device__ unsigned operator()(size_t i) const { return __popc(dir[i].x); }
};
// learn/base/localizer.h:139-166: the ascending unique list, here read off the directory.
struct DirEmit {
  uint2* dir;
  uint64_t* uniq;
  __device__ void operator()(size_t i, unsigned v, unsigned prefix) const {
    dir[i].y = prefix;
    if (v) {
      unsigned bits = dir[i].x;
      uint64_t base = (uint64_t)i << 5;
      while (bits) {
        int b = __ffs(bits) - 1;
        uniq[prefix++] = base + b;
        bits &= bits - 1;
      }
    }
  }
};
// the same, and Handle::Pull for every key it emits (batch.cuh: PullAtEmit)
struct DirEmitPull {
  uint2* dir;
  uint64_t* uniq;
  PullAtEmit p;
  __device__ void operator()(size_t i, unsigned v, unsigned prefix) const {
    dir[i].y = prefix;
    if (!v) return;
    unsigned bits = dir[i].x;
    const uint64_t base = (uint64_t)i << 5;
    // first the loads of all keys of the word (independent), then the stores
    float wv[4];
    unsigned done = 0;
    while (bits) {
      unsigned chunk = 0, bb = bits;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (!bb) break;
        const uint64_t k = base + (__ffs(bb) - 1);
        bb &= bb - 1;
        float w = 0.f;
        if (k >= p.key_begin && k < p.key_end) {
          w = __ldg(&static_cast<const ShardEntry*>(p.table)[k - p.key_begin].w);
        } else {
          atomicAdd(reinterpret_cast<unsigned long long*>(p.counters + 1), 1ull);
        }
        wv[u] = w;
        ++chunk;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (u >= (int)chunk) break;
        uniq[prefix + done + u] = base + (__ffs(bits) - 1);
        p.w[prefix + done + u] = wv[u];
        bits &= bits - 1;
      }
      done += chunk;
    }
  }
};

// The directory scan in one launch (scan.cuh: decoupled look-back). The state buffer belongs to the context.
static int dir_scan(whContext* ctx, uint2* dir, uint64_t* d_uniq, size_t words, unsigned* d_nuniq) {
  static int coop = -1;   // WH_SCAN_COOP=0: the look-back scan for every size
  if (coop < 0) { const char* e = getenv("WH_SCAN_COOP"); coop = e ? atoi(e) : 1; }
  if (coop && scan_coop_fits(ctx, words)) {
    WH_TRY(ctx->scan_tmp.reserve(div_up(words, kScanThreads * kOnepassItems) * sizeof(unsigned) + 256));
    return scan_coop(ctx, DirPopc{dir}, DirEmit{dir, d_uniq}, words, ctx->scan_tmp.as<unsigned>(), d_nuniq);
  }
  const size_t need = scan_onepass_bytes(words);
  if (ctx->scan_state.cap < need) {
    WH_TRY(ctx->scan_state.reserve(need));
    WH_CUDA(cudaMemsetAsync(ctx->scan_state.p, 0, ctx->scan_state.cap, ctx->stream));
    ctx->scan_epoch = 0;
  }
  return scan_onepass(ctx, DirPopc{dir}, DirEmit{dir, d_uniq}, words, ctx->scan_state.p, &ctx->scan_epoch, d_nuniq);
}

// learn/base/localizer.h:179-194,210: local id = rank of the transformed key.
__global__ void __launch_bounds__(256) loc_rank_kernel(const uint64_t* __restrict__ index, size_t n,
                                                       KeyXform xf, const uint2* __restrict__ dir,
                                                       uint32_t* __restrict__ out, uint32_t* cnt) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += stride * kLocIlp) {
    uint64_t k[kLocIlp];
    uint2 e[kLocIlp];
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) { const size_t i = i0 + u * stride; k[u] = i < n ? ld_stream_u64(index + i) : 0; }
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) { k[u] = xf(k[u]); e[u] = __ldg(dir + (k[u] >> 5)); }
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) {
      const size_t i = i0 + u * stride;
      if (i < n) {
        const unsigned r = e[u].y + __popc(e[u].x & ((1u << (k[u] & 31)) - 1u));
        out[i] = r;
        if (cnt) atomicAdd(cnt + r, 1u);
      }
    }
  }
}


// ---- the same two kernels, fed by TMA ------------------------------------------------------------
// Tiles of 2048 ids (16 KB) stream into a 4-deep ring of shared-memory stages by cp.async.bulk; the
// block's threads take 8 ids each from the stage that has landed. Requires a 16-byte aligned id
// array (the plain kernels above remain for the unaligned case).
constexpr int kTileIds = 2048, kTmaStages = 2, kTmaThreads = 256, kPerThread = kTileIds / kTmaThreads;

struct IdRing {
  uint64_t* stage;   // kTmaStages * kTileIds ids
  uint64_t* bar;     // kTmaStages mbarriers
  const uint64_t* index;
  size_t n, ntiles;
  __device__ unsigned tile_ids(size_t t) const { return (unsigned)min((size_t)kTileIds, n - t * kTileIds); }
  __device__ void issue(size_t t, int s) const {   // one thread
    const unsigned bytes = (tile_ids(t) * 8u + 15u) & ~15u;   // the array is padded to 16 B by the caller's contract
    fence_proxy_async();
    mbar_expect_tx(bar + s, bytes);
    bulk_g2s(stage + (size_t)s * kTileIds, index + t * kTileIds, bytes, bar + s);
  }
};

template <typename F>
__device__ __forceinline__ void id_ring_run(const uint64_t* index, size_t n, uint64_t* stage, uint64_t* bar, F body) {
  IdRing r{stage, bar, index, n, (n + kTileIds - 1) / kTileIds};
  if (threadIdx.x == 0) {
    for (int s = 0; s < kTmaStages; ++s) mbar_init(bar + s, 1);
    mbar_init_fence();
  }
  __syncthreads();
  if (threadIdx.x == 0)
    for (int s = 0; s < kTmaStages; ++s) {
      const size_t t = blockIdx.x + (size_t)s * gridDim.x;
      if (t < r.ntiles) r.issue(t, s);
    }
  size_t k = 0;
  for (size_t t = blockIdx.x; t < r.ntiles; t += gridDim.x, ++k) {
    const int s = (int)(k % kTmaStages);
    mbar_wait(bar + s, (unsigned)((k / kTmaStages) & 1));
    body(stage + (size_t)s * kTileIds, t * kTileIds, r.tile_ids(t));
    __syncthreads();   // every thread is done with stage s
    const size_t nt = t + (size_t)kTmaStages * gridDim.x;
    if (threadIdx.x == 0 && nt < r.ntiles) r.issue(nt, s);
  }
}

__global__ void __launch_bounds__(kTmaThreads) loc_mark_tma_kernel(const uint64_t* __restrict__ index, size_t n,
                                                                   KeyXform xf, uint2* dir) {
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* stage = reinterpret_cast<uint64_t*>(smem);
  uint64_t* bar = stage + (size_t)kTmaStages * kTileIds;
  unsigned* cache = reinterpret_cast<unsigned*>(bar + kTmaStages);
  for (int i = threadIdx.x; i < kMarkCache; i += blockDim.x) cache[i] = 0xffffffffu;
  id_ring_run(index, n, stage, bar, [&](const uint64_t* ids, size_t, unsigned cnt) {
    uint64_t k[kPerThread];
    unsigned word[kPerThread];
    bool probe[kPerThread];
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) {
      const unsigned j = u * kTmaThreads + threadIdx.x;
      probe[u] = j < cnt;
      k[u] = xf(probe[u] ? ids[j] : 0);
      const unsigned k32 = (unsigned)k[u];
      const unsigned slot = (k32 * 0x9E3779B1u) >> (32 - 13);
      if (probe[u]) {
        if (cache[slot] == k32) probe[u] = false;
        else cache[slot] = k32;
      }
    }
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) word[u] = probe[u] ? __ldcg(&dir[k[u] >> 5].x) : 0xffffffffu;
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) {
      const unsigned bit = 1u << (k[u] & 31);
      if (probe[u] && !(word[u] & bit)) atomicOr(&dir[k[u] >> 5].x, bit);
    }
  });
}

__global__ void __launch_bounds__(kTmaThreads) loc_rank_tma_kernel(const uint64_t* __restrict__ index, size_t n,
                                                                   KeyXform xf, const uint2* __restrict__ dir,
                                                                   uint32_t* __restrict__ out, uint32_t* cnt_out) {
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* stage = reinterpret_cast<uint64_t*>(smem);
  uint64_t* bar = stage + (size_t)kTmaStages * kTileIds;
  id_ring_run(index, n, stage, bar, [&](const uint64_t* ids, size_t base, unsigned cnt) {
    uint64_t k[kPerThread];
    uint2 e[kPerThread];
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) {
      const unsigned j = u * kTmaThreads + threadIdx.x;
      k[u] = xf(j < cnt ? ids[j] : 0);
      e[u] = __ldg(dir + (k[u] >> 5));
    }
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) {
      const unsigned j = u * kTmaThreads + threadIdx.x;
      if (j < cnt) {
        const unsigned r = e[u].y + __popc(e[u].x & ((1u << (k[u] & 31)) - 1u));
        out[base + j] = r;
        if (cnt_out) atomicAdd(cnt_out + r, 1u);
      }
    }
  });
}

int localize_direct(whContext* ctx, const LocInput& in, uint64_t* d_uniq, uint32_t* d_cnt,
                    uint32_t* d_index, unsigned* d_nuniq) {
  const size_t words = div_up(in.xf.max_key, 32);
  // the directory is kept all-zero between calls; (re)allocation zeroes it once
  if (ctx->directory.cap < words * sizeof(uint2) || ctx->directory_words < words) {
    WH_TRY(ctx->directory.reserve(words * sizeof(uint2)));
    WH_CUDA(cudaMemsetAsync(ctx->directory.p, 0, ctx->directory.cap, ctx->stream));
    ctx->directory_words = ctx->directory.cap / sizeof(uint2);
  }
  uint2* dir = ctx->directory.as<uint2>();
  WH_TRY(ctx->scan_tmp.reserve(div_up(words, kScanTile) * sizeof(unsigned) + 256));
  const size_t n = in.nnz;
  const unsigned grid = (unsigned)std::min<size_t>(div_up(n ? n : 1, 256), (size_t)ctx->num_sms * 16);
  // TMA path: 16-byte aligned ids whose last tile may be read up to the next 16-byte boundary
  // (an even entry count, or at least one spare id of readable memory behind an odd one)
  const bool tma = n >= 2 && ((uintptr_t)in.index & 15) == 0 && (n & 1) == 0;
  const size_t ring_smem = (size_t)kTmaStages * kTileIds * 8 + kTmaStages * 8;
  // (per device, so not cached in a static: a process may hold contexts on several GPUs)
  WH_CUDA(cudaFuncSetAttribute(loc_mark_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(ring_smem + kMarkCache * 4)));
  WH_CUDA(cudaFuncSetAttribute(loc_rank_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ring_smem));
  const unsigned tgrid = (unsigned)std::min<size_t>(div_up(n ? n : 1, kTileIds), (size_t)ctx->num_sms * 2);
  if (n && tma) {
    WH_LAUNCH(loc_mark_tma_kernel, tgrid, kTmaThreads, ring_smem + kMarkCache * 4, ctx->stream, in.index, n, in.xf, dir);
  } else if (n) {
    // Warm-up slice: a small launch sets the presence bits of the features that occur in most
    // rows, so that the full wave does not fire one atomicOr per thread at the same words.
    const size_t head = std::min<size_t>(n, 65536);
    WH_LAUNCH(loc_mark_kernel, (unsigned)div_up(head, 1024), 256, 0, ctx->stream, in.index, head, in.xf, dir);
    if (n > head) {
      const unsigned g2 = (unsigned)std::min<size_t>(div_up(n - head, 256), (size_t)ctx->num_sms * 6);
      WH_LAUNCH(loc_mark_kernel, g2, 256, 0, ctx->stream, in.index + head, n - head, in.xf, dir);
    }
  }
  WH_PROF(ctx, "loc_mark");
  WH_TRY(dir_scan(ctx, dir, d_uniq, words, d_nuniq));
  WH_PROF(ctx, "loc_dir_scan");
  if (d_cnt && n) WH_CUDA(cudaMemsetAsync(d_cnt, 0, std::min<uint64_t>(n, in.xf.max_key) * 4, ctx->stream));
  // (rank stays on the plain kernel: its ids come back from L2, and more resident threads keep more
  //  directory gathers in flight -- measured 20 us vs 24 us for the TMA-fed variant)
  static int rank_tma = -1;
  if (rank_tma < 0) { const char* e = getenv("WH_RANK_TMA"); rank_tma = e ? atoi(e) : 0; }
  if (n && tma && rank_tma) WH_LAUNCH(loc_rank_tma_kernel, tgrid, kTmaThreads, ring_smem, ctx->stream, in.index, n, in.xf, dir, d_index, d_cnt);
  else if (n) WH_LAUNCH(loc_rank_kernel, grid, 256, 0, ctx->stream, in.index, n, in.xf, dir, d_index, d_cnt);
  WH_PROF(ctx, "loc_rank");
  WH_CUDA(cudaMemsetAsync(dir, 0, words * sizeof(uint2), ctx->stream));
  WH_PROF(ctx, "loc_dir_clear");
  return WH_OK;
}

// ------------------------------------------------------------------ direct, with hot keys --
// See hot.cuh. Same directory, same scan, same ascending key list; entries of table keys never touch the
// directory (one shared-memory probe instead of an L2 access in mark and in rank) and come out as codes.

constexpr int kHotMarkThreads = kTmaThreads;
// Entries that are not table keys leave the mark pass holding their transformed key (< 2^28, flag bit clear); the
// rank pass turns it into the rank in place, so that it never has to read the 64-bit ids again.
__global__ void __launch_bounds__(kTmaThreads) loc_mark_hot_kernel(const uint64_t* __restrict__ index, size_t n,
                                                                   KeyXform xf, uint2* dir, HotView hot,
                                                                   uint32_t* __restrict__ codes) {
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* stage = reinterpret_cast<uint64_t*>(smem);
  uint64_t* bar = stage + (size_t)kTmaStages * kTileIds;
  static_assert(kTmaStages % 2 == 0, "the tag table behind the barriers must stay 16-byte aligned");
  uint4* s_tags = reinterpret_cast<uint4*>(bar + kTmaStages);
  unsigned char* s_present = reinterpret_cast<unsigned char*>(s_tags + kHotBuckets);
  hot_load_tags(s_tags, hot.tags);
  for (int i = threadIdx.x; i < kHotSlots / 4; i += blockDim.x) reinterpret_cast<unsigned*>(s_present)[i] = 0u;
  // (id_ring_run starts with a __syncthreads)
  id_ring_run(index, n, stage, bar, [&](const uint64_t* ids, size_t base, unsigned cnt) {
    uint64_t k[kPerThread];
    unsigned word[kPerThread];
    bool probe[kPerThread];
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) {
      const unsigned j = u * kTmaThreads + threadIdx.x;
      probe[u] = j < cnt;
      k[u] = xf(probe[u] ? ids[j] : 0);
      if (probe[u]) {
        const int slot = hot_probe(s_tags, (unsigned)k[u]);
        codes[base + j] = slot >= 0 ? (kHotFlag | (unsigned)slot) : (unsigned)k[u];   // code, or the key for the rank pass
        if (slot >= 0) { s_present[slot] = 1; probe[u] = false; }   // same-value byte stores: no atomics needed
      }
    }
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) word[u] = probe[u] ? __ldcg(&dir[k[u] >> 5].x) : 0xffffffffu;
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) {
      const unsigned bit = 1u << (k[u] & 31);
      if (probe[u] && !(word[u] & bit)) atomicOr(&dir[k[u] >> 5].x, bit);
    }
  });
  __syncthreads();
  for (int i = threadIdx.x; i < kHotSlots; i += blockDim.x) if (s_present[i]) hot.present[i] = 1u;
}

// The same pass without the TMA ring: a plain grid-stride kernel whose only shared memory is the table (40 KB), so
// that four times as many warps are resident to hide the directory probes of the entries that miss it.
constexpr int kHotMarkPlainThreads = 512;
// POW2: max_key is a power of two (the usual hashed key space), so the transform is a mask and everything after
// the id load is 32-bit; the general instantiation keeps KeyXform's run-time switch (modulo / nibble reversal).
// n < 2^31 (checked at the launch), so positions are 32-bit too. Measured effect of the specialisation: 33.6 -> 32.8 us.
template <bool POW2>
__global__ void __launch_bounds__(kHotMarkPlainThreads) loc_mark_hot_plain_kernel(const uint64_t* __restrict__ index, unsigned n,
                                                                                  KeyXform xf, uint2* dir, HotView hot,
                                                                                  uint32_t* __restrict__ codes) {
  __shared__ uint4 s_tags[kHotBuckets];
  __shared__ unsigned char s_present[kHotSlots];
  hot_load_tags(s_tags, hot.tags);
  for (int i = threadIdx.x; i < kHotSlots / 4; i += blockDim.x) reinterpret_cast<unsigned*>(s_present)[i] = 0u;
  __syncthreads();
  const unsigned mask = (unsigned)(xf.max_key - 1);
  const unsigned stride = gridDim.x * blockDim.x;
  // i0 + (kLocIlp - 1) * stride must not wrap: the launch keeps gridDim * blockDim * kLocIlp <= 2^31
  for (unsigned i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += stride * kLocIlp) {
    uint64_t id[kLocIlp];
    unsigned k[kLocIlp], word[kLocIlp];
    bool probe[kLocIlp];
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) {
      const unsigned i = i0 + u * stride;
      probe[u] = i < n;
      id[u] = probe[u] ? ld_stream_u64(index + i) : 0;
    }
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) {
      k[u] = POW2 ? ((unsigned)id[u] & mask) : (unsigned)xf(id[u]);     // direct addressing: keys < 2^28
      if (probe[u]) {
        const int slot = hot_probe(s_tags, k[u]);
        codes[i0 + u * stride] = slot >= 0 ? (kHotFlag | (unsigned)slot) : k[u];   // code, or the key for the rank pass
        if (slot >= 0) { s_present[slot] = 1; probe[u] = false; }
      }
    }
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) word[u] = probe[u] ? __ldcg(&dir[k[u] >> 5].x) : 0xffffffffu;
#pragma unroll
    for (int u = 0; u < kLocIlp; ++u) {
      const unsigned bit = 1u << (k[u] & 31);
      if (!(word[u] & bit)) atomicOr(&dir[k[u] >> 5].x, bit);   // (word is all ones where nothing is to be done)
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kHotSlots; i += blockDim.x) if (s_present[i]) hot.present[i] = 1u;
}

// table keys that occur in this minibatch enter the directory like any other key (kHotSlots threads; folding
// this into the mark kernel's tail -- last block, or per-block test-and-set -- cost 15-25 us of serial latency)
__global__ void hot_present_kernel(HotView hot, uint2* dir) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= kHotSlots || !hot.present[p]) return;
  const unsigned k = reinterpret_cast<const unsigned*>(hot.tags)[p];
  atomicOr(&dir[k >> 5].x, 1u << (k & 31));
}
// after the scan: the rank of every table key that is present; resets the presence flags for the next minibatch
__device__ __forceinline__ void hot_rank_slot(const HotView& hot, const uint2* __restrict__ dir, int p) {
  unsigned r = kHotNone;
  if (hot.present[p]) {
    const unsigned k = reinterpret_cast<const unsigned*>(hot.tags)[p];
    const uint2 e = dir[k >> 5];
    r = e.y + __popc(e.x & ((1u << (k & 31)) - 1u));
    hot.present[p] = 0u;
  }
  hot.rank[p] = r;
}

// the table keys' ranks as the scan's epilogue (a slice per block), or as a launch of their own after the look-back scan
struct HotRankEpilogue {
  HotView hot;
  uint2* dir;
  static constexpr bool kPresent = true;
  // before the scan: table keys that occur in this minibatch enter the directory like any other key
  __device__ void before() const {
    const int per = (kHotSlots + (int)gridDim.x - 1) / (int)gridDim.x;
    const int lo = (int)blockIdx.x * per, hi = min(kHotSlots, lo + per);
    for (int p = lo + (int)threadIdx.x; p < hi; p += (int)blockDim.x) {
      if (__ldcg(hot.present + p)) {
        const unsigned k = reinterpret_cast<const unsigned*>(hot.tags)[p];
        atomicOr(&dir[k >> 5].x, 1u << (k & 31));
      }
    }
  }
  __device__ void operator()() const {
    const int per = (kHotSlots + (int)gridDim.x - 1) / (int)gridDim.x;
    const int lo = (int)blockIdx.x * per, hi = min(kHotSlots, lo + per);
    for (int p = lo + (int)threadIdx.x; p < hi; p += (int)blockDim.x) hot_rank_slot(hot, dir, p);
  }
};
__global__ void hot_rank_kernel(HotView hot, const uint2* __restrict__ dir) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < kHotSlots) hot_rank_slot(hot, dir, p);
}
// directory scan that also leaves hot.rank filled (and hot.present reset)
// (also enters the present table keys into the directory first: as the cooperative scan's first phase, or as a launch of
//  its own in front of the look-back scan)
static int dir_scan_hot(whContext* ctx, uint2* dir, uint64_t* d_uniq, size_t words, unsigned* d_nuniq, const HotView& hot,
                        const PullAtEmit* pull, bool* pulled) {
  // One cooperative launch (scan + the table keys' ranks) or five plain ones. Alone on the GPU the cooperative scan is the
  // shorter (20 against 23.5 us); in a sharded job with several minibatches in flight it has to become co-resident beside
  // the cooperative apply and the forward/backward kernel of its neighbours, and the plain launches win: 0.139 against
  // 0.146 ms per step on 2 GPUs. WH_SCAN_COOP=0|1 forces one.
  static int forced = -2;
  if (forced == -2) { const char* e = getenv("WH_SCAN_COOP"); forced = e ? atoi(e) : -1; }
  const bool coop = forced >= 0 ? forced != 0 : !ctx->avoid_coop;
  *pulled = false;
  if (coop && scan_coop_fits(ctx, words)) {
    WH_TRY(ctx->scan_tmp.reserve(div_up(words, kScanThreads * kOnepassItems) * sizeof(unsigned) + 256));
    if (pull && pull->table) {
      if (pull->after) WH_CUDA(cudaStreamWaitEvent(ctx->stream, pull->after, 0));
      *pulled = true;
      return scan_coop(ctx, DirPopc{dir}, DirEmitPull{dir, d_uniq, *pull}, words, ctx->scan_tmp.as<unsigned>(), d_nuniq,
                       HotRankEpilogue{hot, dir});
    }
    return scan_coop(ctx, DirPopc{dir}, DirEmit{dir, d_uniq}, words, ctx->scan_tmp.as<unsigned>(), d_nuniq, HotRankEpilogue{hot, dir});
  }
  WH_LAUNCH(hot_present_kernel, kHotSlots / 256, 256, 0, ctx->stream, hot, dir);
  WH_TRY(dir_scan(ctx, dir, d_uniq, words, d_nuniq));
  WH_LAUNCH(hot_rank_kernel, kHotSlots / 256, 256, 0, ctx->stream, hot, dir);
  return WH_OK;
}

// rank pass: entries that are not table keys hold their key; the directory turns it into the rank
__global__ void __launch_bounds__(256) loc_rank_cold_kernel(size_t n, const uint2* __restrict__ dir, uint32_t* __restrict__ codes) {
  constexpr int kIlp = 8;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += stride * kIlp) {
    uint32_t c[kIlp];
    uint2 e[kIlp];
#pragma unroll
    for (int u = 0; u < kIlp; ++u) { const size_t i = i0 + u * stride; c[u] = i < n ? ld_stream_u32(codes + i) : kHotFlag; }
#pragma unroll
    for (int u = 0; u < kIlp; ++u) if (!(c[u] & kHotFlag)) e[u] = __ldg(dir + (c[u] >> 5));
#pragma unroll
    for (int u = 0; u < kIlp; ++u)
      if (!(c[u] & kHotFlag)) codes[i0 + u * stride] = e[u].y + __popc(e[u].x & ((1u << (c[u] & 31)) - 1u));
  }
}

bool pull_at_emit_supported(const whContext* ctx, uint64_t max_key) {
  static int coop = -1;
  if (coop < 0) { const char* e = getenv("WH_SCAN_COOP"); coop = e ? atoi(e) : 1; }
  // Opt-in (WH_PULL_AT_EMIT=1). Measured at the Criteo shape: one launch and 5 us of kernel time less per minibatch, and
  // 848 vs 830 M examples/s with three minibatches in flight -- but the scan now has to wait for the previous push, which
  // puts it on the push -> pull -> forward/backward -> push cycle, and with four in flight that cycle becomes the limit
  // (860 vs 890 M examples/s). The default keeps the separate pull launch.
  const char* on = getenv("WH_PULL_AT_EMIT");
  return coop && on && atoi(on) == 1 && direct_supported(max_key) && scan_coop_fits(ctx, div_up(max_key, 32));
}

int localize_direct_hot(whContext* ctx, const LocInput& in, const HotView& hot, uint64_t* d_uniq, uint32_t* d_codes,
                        unsigned* d_nuniq, bool resolve, const PullAtEmit* pull) {
  const size_t words = div_up(in.xf.max_key, 32);
  if (ctx->directory.cap < words * sizeof(uint2) || ctx->directory_words < words) {
    WH_TRY(ctx->directory.reserve(words * sizeof(uint2)));
    WH_CUDA(cudaMemsetAsync(ctx->directory.p, 0, ctx->directory.cap, ctx->stream));
    ctx->directory_words = ctx->directory.cap / sizeof(uint2);
  }
  uint2* dir = ctx->directory.as<uint2>();
  WH_TRY(ctx->scan_tmp.reserve(div_up(words, kScanTile) * sizeof(unsigned) + 256));
  const size_t n = in.nnz;
  const bool tma = n >= 2 && ((uintptr_t)in.index & 15) == 0 && (n & 1) == 0;
  const size_t mark_smem = (size_t)kTmaStages * kTileIds * 8 + kTmaStages * 8 + kHotSlots * 4 + kHotSlots;
  WH_CUDA(cudaFuncSetAttribute(loc_mark_hot_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mark_smem));
  const unsigned tgrid = (unsigned)std::min<size_t>(div_up(n, kTileIds), (size_t)ctx->num_sms * 3);
  // measured at the Criteo shape: plain kernel, 2 x 512 threads per SM 33.8 us; TMA ring (2 stages, 3 blocks per SM)
  // 40.1 us -- with 78 % of the entries answered from shared memory the pass is bound by instruction issue, and
  // the ring's extra shared memory only costs resident warps. WH_HOT_MARK_PLAIN=0 selects the ring.
  static int plain = -1;
  if (plain < 0) { const char* e = getenv("WH_HOT_MARK_PLAIN"); plain = e ? atoi(e) : 2; }
  if (plain || !tma) {   // (the ring needs 16-byte aligned ids and an even count)
    const unsigned pg = (unsigned)std::min<size_t>(div_up(n, kHotMarkPlainThreads * kLocIlp), (size_t)ctx->num_sms * (plain ? plain : 2));
    WH_REQUIRE(n < (1ull << 31), "localize_direct_hot: %zu entries", n);
    if (in.xf.mode == 1) WH_LAUNCH(loc_mark_hot_plain_kernel<true>, pg, kHotMarkPlainThreads, 0, ctx->stream, in.index, (unsigned)n, in.xf, dir, hot, d_codes);
    else WH_LAUNCH(loc_mark_hot_plain_kernel<false>, pg, kHotMarkPlainThreads, 0, ctx->stream, in.index, (unsigned)n, in.xf, dir, hot, d_codes);
  } else {
    WH_LAUNCH(loc_mark_hot_kernel, tgrid, kTmaThreads, mark_smem, ctx->stream, in.index, n, in.xf, dir, hot, d_codes);
  }
  WH_PROF(ctx, "loc_mark");
  bool pulled = false;
  WH_TRY(dir_scan_hot(ctx, dir, d_uniq, words, d_nuniq, hot, pull, &pulled));
  WH_REQUIRE(!(pull && pull->table) || pulled, "localize_direct_hot: the pull could not ride in the scan (directory too large for the cooperative scan)");
  WH_PROF(ctx, "loc_dir_scan");
  if (resolve) {
    const unsigned rgrid = (unsigned)std::min<size_t>(div_up(n, 256 * 8), (size_t)ctx->num_sms * 8);
    WH_LAUNCH(loc_rank_cold_kernel, rgrid, 256, 0, ctx->stream, n, dir, d_codes);
    WH_PROF(ctx, "loc_rank");
    WH_CUDA(cudaMemsetAsync(dir, 0, words * sizeof(uint2), ctx->stream));
    WH_PROF(ctx, "loc_dir_clear");
  }
  // (!resolve: entries that are not table keys still hold their key, and the directory stays as the scan left it --
  //  the caller turns keys into ranks where it reads the codes, and zeroes the directory afterwards)
  return WH_OK;
}

// Most frequent keys first; a key whose bucket is already full is left out (it simply stays an ordinary key).
int hot_build_image(const uint64_t* keys, const uint32_t* counts, size_t n, uint32_t* image) {
  std::fill(image, image + kHotSlots, kHotEmpty);
  std::vector<uint32_t> order(n);
  std::iota(order.begin(), order.end(), 0u);
  const size_t want = std::min<size_t>(n, kHotSlots / 2);          // half-full table: few buckets overflow
  std::partial_sort(order.begin(), order.begin() + want, order.end(), [&](uint32_t a, uint32_t b) {
    return counts[a] != counts[b] ? counts[a] > counts[b] : keys[a] < keys[b];
  });
  int placed = 0;
  for (size_t i = 0; i < want; ++i) {
    const uint64_t k = keys[order[i]];
    if (counts[order[i]] < 2 || k >= kDirectMaxKey) continue;
    const unsigned k32 = (unsigned)k;
    const unsigned b = (k32 * 0x9E3779B1u) >> (32 - 11);
    for (int w = 0; w < kHotWays; ++w)
      if (image[b * kHotWays + w] == kHotEmpty) { image[b * kHotWays + w] = k32; ++placed; break; }
  }
  return placed;
}

// --------------------------------------------------------------------------------- sort --

// learn/base/localizer.h:107-128: pair[i] = {T(index[i]), i}; also folds the OR / AND of all
// keys so that the sort can skip constant digits.
__global__ void __launch_bounds__(256) loc_xform_kernel(const uint64_t* __restrict__ index, size_t n,
                                                        KeyXform xf, uint64_t* keys, uint32_t* vals,
                                                        SortPlan* plan) {
  unsigned long long o = 0ull, a = ~0ull;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t k = xf(ld_stream_u64(index + i));
    keys[i] = k;
    vals[i] = (uint32_t)i;
    o |= k;
    a &= k;
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    o |= __shfl_xor_sync(0xffffffffu, o, s);
    a &= __shfl_xor_sync(0xffffffffu, a, s);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicOr(&plan->bits_or, o);
    atomicAnd(&plan->bits_and, a);
  }
}

// learn/base/localizer.h:139-166 (run-length encode the sorted pairs) and :179-194 (rank -> position)
struct HeadFlag {
  SortBufs b;
  __device__ unsigned operator()(size_t i) const {
    const uint64_t* k = sorted_keys(b);
    return (i == 0 || k[i] != k[i - 1]) ? 1u : 0u;
  }
};
struct HeadApply {
  SortBufs b;
  uint64_t* uniq;
  uint32_t* headpos;
  uint32_t* index;
  __device__ void operator()(size_t i, unsigned v, unsigned prefix) const {
    unsigned rank = prefix + v - 1;
    index[sorted_vals(b)[i]] = rank;
    if (v) { uniq[rank] = sorted_keys(b)[i]; headpos[rank] = (uint32_t)i; }
  }
};
__global__ void loc_count_kernel(const uint32_t* __restrict__ headpos, const unsigned* nuniq, size_t n,
                                 uint32_t* cnt) {
  const unsigned U = *nuniq;
  for (size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x; r < U; r += (size_t)gridDim.x * blockDim.x)
    cnt[r] = (r + 1 < U ? headpos[r + 1] : (uint32_t)n) - headpos[r];
}

int localize_sort(whContext* ctx, const LocInput& in, uint64_t* d_uniq, uint32_t* d_cnt,
                  uint32_t* d_index, unsigned* d_nuniq) {
  const size_t n = in.nnz;
  if (n == 0) { WH_CUDA(cudaMemsetAsync(d_nuniq, 0, sizeof(unsigned), ctx->stream)); return WH_OK; }
  SortBufs b;
  WH_TRY(ctx->sort_tmp[0].reserve(n * 8)); WH_TRY(ctx->sort_tmp[1].reserve(n * 8));
  WH_TRY(ctx->sort_tmp[2].reserve(n * 4)); WH_TRY(ctx->sort_tmp[3].reserve(n * 4));
  WH_TRY(ctx->sort_tmp[4].reserve(sort_counts_bytes(n) + sizeof(SortPlan) + 256));
  WH_TRY(ctx->sort_tmp[5].reserve(n * 4));
  WH_TRY(ctx->scan_tmp.reserve(div_up(n, kScanTile) * sizeof(unsigned) + 256));
  b.keys[0] = ctx->sort_tmp[0].as<uint64_t>(); b.keys[1] = ctx->sort_tmp[1].as<uint64_t>();
  b.vals[0] = ctx->sort_tmp[2].as<uint32_t>(); b.vals[1] = ctx->sort_tmp[3].as<uint32_t>();
  b.plan = ctx->sort_tmp[4].as<SortPlan>();
  b.counts = reinterpret_cast<uint32_t*>(ctx->sort_tmp[4].as<char>() + 256);
  uint32_t* headpos = ctx->sort_tmp[5].as<uint32_t>();
  WH_TRY(sort_plan_reset(ctx, b.plan));
  const unsigned grid = (unsigned)std::min<size_t>(div_up(n, 256), (size_t)ctx->num_sms * 16);
  WH_LAUNCH(loc_xform_kernel, grid, 256, 0, ctx->stream, in.index, n, in.xf, b.keys[0], b.vals[0], b.plan);
  WH_PROF(ctx, "loc_xform");
  WH_TRY(radix_sort_pairs(ctx, b, n));
  WH_PROF(ctx, "loc_radix_sort");
  WH_TRY(scan_and_apply<unsigned>(ctx, HeadFlag{b}, HeadApply{b, d_uniq, headpos, d_index}, n,
                        ctx->scan_tmp.as<unsigned>(), d_nuniq));
  if (d_cnt) WH_LAUNCH(loc_count_kernel, grid, 256, 0, ctx->stream, headpos, d_nuniq, n, d_cnt);
  WH_PROF(ctx, "loc_heads_rank");
  return WH_OK;
}

// --------------------------------------------------------- RemapIndex with a foreign dict --

// dictpos[u] = 1 + position of uniq[u] in the ascending dictionary, 0 if absent
// (the two-pointer merge of learn/base/localizer.h:179-194, as a search per unique key).
__global__ void loc_dictpos_kernel(const uint64_t* __restrict__ uniq, const unsigned* nuniq,
                                   const uint64_t* __restrict__ dict, size_t ndict, uint32_t* dictpos) {
  const unsigned U = *nuniq;
  for (size_t u = (size_t)blockIdx.x * blockDim.x + threadIdx.x; u < U; u += (size_t)gridDim.x * blockDim.x) {
    uint64_t k = uniq[u];
    size_t lo = 0, hi = ndict;
    while (lo < hi) { size_t mid = (lo + hi) >> 1; if (dict[mid] < k) lo = mid + 1; else hi = mid; }
    dictpos[u] = (lo < ndict && dict[lo] == k) ? (uint32_t)(lo + 1) : 0u;
  }
}
struct KeepFlag {
  const uint32_t* index; const uint32_t* dictpos;
  __device__ unsigned operator()(size_t i) const { return dictpos[index[i]] != 0 ? 1u : 0u; }
};
// learn/base/localizer.h:203-213: drop unmatched entries, keep order.
struct KeepApply {
  const uint32_t* index; const uint32_t* dictpos; const float* value;
  uint32_t* scanpos; uint32_t* out_index; float* out_value;
  __device__ void operator()(size_t i, unsigned v, unsigned prefix) const {
    scanpos[i] = prefix;
    if (v) { out_index[prefix] = dictpos[index[i]] - 1; if (value) out_value[prefix] = value[i]; }
  }
};
__global__ void loc_reoffset_kernel(const size_t* __restrict__ offset, size_t rows, size_t n,
                                    const uint32_t* __restrict__ scanpos, const unsigned* total,
                                    size_t* out) {
  for (size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x; r <= rows; r += (size_t)gridDim.x * blockDim.x) {
    size_t o = offset[r];
    out[r] = o < n ? scanpos[o] : *total;
  }
}

}  // namespace wh

using namespace wh;

static int localize_common(whContext* ctx, const whRowBlock* blk, int mem, uint64_t max_key, int method,
                           const uint64_t* dict, size_t ndict, bool with_dict, whBatch** out) {
  WH_REQUIRE(ctx && blk && out, "wh_localize: NULL argument");
  WH_REQUIRE(mem == WH_HOST || mem == WH_DEVICE, "wh_localize: bad mem %d", mem);
  WH_REQUIRE(max_key != 0, "wh_localize: max_key must be positive");
  DeviceGuard g(ctx->device);
  const bool want_cnt = (method & 16) != 0;
  method &= 15;
  whBatch* b = new whBatch();
  b->ctx = ctx;
  *out = b;
  if (blk->size == 0) return WH_OK;  // learn/base/localizer.h:101,173: nothing is written
  WH_REQUIRE(blk->offset != nullptr && (blk->index != nullptr), "wh_localize: offset/index is NULL");
  const size_t rows = blk->size;
  // the entry count lives in offset[rows]
  size_t n = 0, first = 0;
  if (mem == WH_HOST) { n = blk->offset[rows]; first = blk->offset[0]; }
  else {
    WH_CUDA(cudaMemcpyAsync(ctx->pinned.p, blk->offset + rows, sizeof(size_t), cudaMemcpyDeviceToHost, ctx->stream));
    WH_CUDA(cudaMemcpyAsync(ctx->pinned.as<size_t>() + 1, blk->offset, sizeof(size_t), cudaMemcpyDeviceToHost, ctx->stream));
    WH_CUDA(cudaStreamSynchronize(ctx->stream));
    n = ctx->pinned.as<size_t>()[0]; first = ctx->pinned.as<size_t>()[1];
  }
  WH_REQUIRE(first == 0, "wh_localize: offset[0] must be 0 (learn/base/localizer.h:102)");
  WH_REQUIRE(n < (size_t)UINT32_MAX, "wh_localize: %zu entries; positions are 32-bit (learn/base/localizer.h:103)", n);
  if (method == WH_LOC_AUTO) method = direct_supported(max_key) ? WH_LOC_DIRECT : WH_LOC_SORT;
  WH_REQUIRE(method == WH_LOC_SORT || direct_supported(max_key),
             "wh_localize: direct addressing needs max_key <= 2^28");

  const uint64_t* d_index; const size_t* d_offset; const float* d_value; const float* d_label;
  WH_TRY(to_device(ctx, ctx->stage[0], blk->index, n, mem, &d_index));
  // offset / value / label are copied into the batch (Localize hands back its own container)
  b->rows = rows;
  WH_TRY(b->offset.reserve((rows + 1) * sizeof(size_t)));
  cudaMemcpyKind kind = mem == WH_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
  WH_CUDA(cudaMemcpyAsync(b->offset.p, blk->offset, (rows + 1) * sizeof(size_t), kind, ctx->stream));
  d_offset = b->offset.as<size_t>();
  if (blk->label) {
    WH_TRY(b->label.reserve(rows * 4));
    WH_CUDA(cudaMemcpyAsync(b->label.p, blk->label, rows * 4, kind, ctx->stream));
    b->has_label = true;
  }
  d_label = b->label.as<float>(); (void)d_label;
  d_value = nullptr;
  if (blk->value && n) {
    WH_TRY(b->value.reserve(n * 4));
    WH_CUDA(cudaMemcpyAsync(b->value.p, blk->value, n * 4, kind, ctx->stream));
    b->has_value = true;
    d_value = b->value.as<float>();
  }
  const size_t ucap = (size_t)std::min<uint64_t>(n ? n : 1, max_key);
  WH_TRY(b->uniq.reserve(ucap * 8));
  WH_TRY(b->index.reserve((n ? n : 1) * 4));
  if (want_cnt) { WH_TRY(b->cnt.reserve(ucap * 4)); b->has_cnt = true; }
  unsigned* d_nuniq = ctx->small.as<unsigned>();
  LocInput in{d_index, n, make_xform(max_key)};
  if (method == WH_LOC_DIRECT)
    WH_TRY(localize_direct(ctx, in, b->uniq.as<uint64_t>(), want_cnt ? b->cnt.as<uint32_t>() : nullptr,
                           b->index.as<uint32_t>(), d_nuniq));
  else
    WH_TRY(localize_sort(ctx, in, b->uniq.as<uint64_t>(), want_cnt ? b->cnt.as<uint32_t>() : nullptr,
                         b->index.as<uint32_t>(), d_nuniq));
  b->nnz = n;

  unsigned* d_total = d_nuniq + 1;
  if (with_dict) {
    if (ndict == 0 || n == 0) {
      // RemapIndex returns before touching `localized` (learn/base/localizer.h:173)
      b->rows = 0; b->nnz = 0; b->has_value = b->has_label = false;
    } else {
      const uint64_t* d_dict;
      WH_TRY(to_device(ctx, ctx->stage[1], dict, ndict, mem, &d_dict));
      WH_TRY(ctx->stage[2].reserve(ucap * 4));           // dictpos
      WH_TRY(ctx->stage[3].reserve(n * 4));              // scanpos
      WH_TRY(ctx->stage[4].reserve(n * 4));              // compacted index
      WH_TRY(ctx->stage[5].reserve(n * 4));              // compacted value
      WH_TRY(ctx->stage[6].reserve((rows + 1) * sizeof(size_t)));
      WH_TRY(ctx->scan_tmp.reserve(div_up(n, kScanTile) * sizeof(unsigned) + 256));
      uint32_t* dictpos = ctx->stage[2].as<uint32_t>();
      const unsigned grid = (unsigned)std::min<size_t>(div_up(ucap, 256), (size_t)ctx->num_sms * 8);
      WH_LAUNCH(loc_dictpos_kernel, grid, 256, 0, ctx->stream, b->uniq.as<uint64_t>(), d_nuniq, d_dict, ndict, dictpos);
      KeepFlag kf{b->index.as<uint32_t>(), dictpos};
      KeepApply ka{b->index.as<uint32_t>(), dictpos, d_value, ctx->stage[3].as<uint32_t>(),
                   ctx->stage[4].as<uint32_t>(), ctx->stage[5].as<float>()};
      WH_TRY(scan_and_apply<unsigned>(ctx, kf, ka, n, ctx->scan_tmp.as<unsigned>(), d_total));
      const unsigned g2 = (unsigned)std::min<size_t>(div_up(rows + 1, 256), (size_t)ctx->num_sms * 8);
      WH_LAUNCH(loc_reoffset_kernel, g2, 256, 0, ctx->stream, d_offset, rows, n, ctx->stage[3].as<uint32_t>(),
                d_total, ctx->stage[6].as<size_t>());
      WH_CUDA(cudaMemcpyAsync(b->offset.p, ctx->stage[6].p, (rows + 1) * sizeof(size_t), cudaMemcpyDeviceToDevice, ctx->stream));
      WH_CUDA(cudaMemcpyAsync(b->index.p, ctx->stage[4].p, n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
      if (d_value) WH_CUDA(cudaMemcpyAsync(b->value.p, ctx->stage[5].p, n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    }
  }
  WH_CUDA(cudaMemcpyAsync(ctx->pinned.p, d_nuniq, 2 * sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  b->nuniq = ctx->pinned.as<unsigned>()[0];
  if (with_dict && b->rows) {
    b->nnz = ctx->pinned.as<unsigned>()[1];
    if (b->nnz == 0) b->has_value = false;
  }
  return WH_OK;
}

extern "C" {

int wh_localize(whContext* ctx, const whRowBlock* blk, int mem, uint64_t max_key, int method, whBatch** out) {
  int r = localize_common(ctx, blk, mem, max_key, method, nullptr, 0, false, out);
  if (r != WH_OK && out && *out) { delete *out; *out = nullptr; }
  return r;
}
int wh_localize_with_dict(whContext* ctx, const whRowBlock* blk, int mem, uint64_t max_key,
                          const uint64_t* dict, size_t ndict, whBatch** out) {
  int r = localize_common(ctx, blk, mem, max_key, WH_LOC_AUTO | 16, dict, ndict, true, out);
  if (r != WH_OK && out && *out) { delete *out; *out = nullptr; }
  return r;
}
size_t wh_batch_rows(const whBatch* b) { return b ? b->rows : 0; }
size_t wh_batch_nnz(const whBatch* b) { return b ? b->nnz : 0; }
size_t wh_batch_num_uniq(const whBatch* b) { return b ? b->nuniq : 0; }
int wh_batch_has_value(const whBatch* b) { return b && b->has_value; }
int wh_batch_has_label(const whBatch* b) { return b && b->has_label; }

__global__ void cnt_to_float_kernel(const uint32_t* c, size_t n, float* out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (float)c[i];  // learn/base/localizer.h:151,163: static_cast<C>(cnt), C = float
}

int wh_batch_copy(whContext* ctx, const whBatch* b, uint64_t* uniq, uint32_t* cnt_u32, float* cnt_f32,
                  size_t* offset, uint32_t* index, float* value, float* label) {
  WH_REQUIRE(ctx && b, "wh_batch_copy: NULL argument");
  DeviceGuard g(ctx->device);
  cudaStream_t s = ctx->stream;
  if (uniq && b->nuniq) WH_CUDA(cudaMemcpyAsync(uniq, b->uniq.p, b->nuniq * 8, cudaMemcpyDeviceToHost, s));
  if ((cnt_u32 || cnt_f32) && b->nuniq) {
    WH_REQUIRE(b->has_cnt, "wh_batch_copy: counts were not requested (method | 16)");
    if (cnt_u32) WH_CUDA(cudaMemcpyAsync(cnt_u32, b->cnt.p, b->nuniq * 4, cudaMemcpyDeviceToHost, s));
    if (cnt_f32) {
      WH_TRY(ctx->stage[7].reserve(b->nuniq * 4));
      WH_LAUNCH(cnt_to_float_kernel, (unsigned)div_up(b->nuniq, 256), 256, 0, s, b->cnt.as<uint32_t>(), b->nuniq, ctx->stage[7].as<float>());
      WH_CUDA(cudaMemcpyAsync(cnt_f32, ctx->stage[7].p, b->nuniq * 4, cudaMemcpyDeviceToHost, s));
    }
  }
  if (offset && b->rows) WH_CUDA(cudaMemcpyAsync(offset, b->offset.p, (b->rows + 1) * sizeof(size_t), cudaMemcpyDeviceToHost, s));
  if (index && b->nnz) WH_CUDA(cudaMemcpyAsync(index, b->index.p, b->nnz * 4, cudaMemcpyDeviceToHost, s));
  if (value && b->nnz && b->has_value) WH_CUDA(cudaMemcpyAsync(value, b->value.p, b->nnz * 4, cudaMemcpyDeviceToHost, s));
  if (label && b->rows && b->has_label) WH_CUDA(cudaMemcpyAsync(label, b->label.p, b->rows * 4, cudaMemcpyDeviceToHost, s));
  WH_CUDA(cudaStreamSynchronize(s));
  return WH_OK;
}

int wh_batch_device_view(const whBatch* b, const uint64_t** uniq, const uint32_t** cnt, whLocalBlock* local) {
  WH_REQUIRE(b, "wh_batch_device_view: NULL batch");
  if (uniq) *uniq = b->uniq.as<uint64_t>();
  if (cnt) *cnt = b->has_cnt ? b->cnt.as<uint32_t>() : nullptr;
  if (local) {
    local->size = b->rows;
    local->offset = b->offset.as<size_t>();
    local->label = b->has_label ? b->label.as<float>() : nullptr;
    local->index = b->index.as<uint32_t>();
    local->value = b->has_value ? b->value.as<float>() : nullptr;
  }
  return WH_OK;
}

int wh_batch_destroy(whBatch* b) {
  if (!b) return WH_OK;
  DeviceGuard g(b->ctx ? b->ctx->device : 0);
  if (b->ctx) cudaStreamSynchronize(b->ctx->stream);
  delete b;
  return WH_OK;
}

}  // extern "C"
