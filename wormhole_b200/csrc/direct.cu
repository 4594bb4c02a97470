// AsgdWorker::ProcessMinibatch (learn/linear/async_sgd.h:240-288) when worker and server share the GPU and the key
// space is hashed (ps::FLAGS_max_key <= 2^28): two launches per minibatch.
//
// The reference localizes the minibatch (learn/base/localizer.h:96-221) because its pull and push are MESSAGES: the
// unique key list is what travels, and the local ids index the values that come back. With the shard in the same HBM
// nothing travels, so nothing is localized: a transformed key addresses its entry of the shard directly -- weight and
// gradient accumulator sit in the same 32-byte sector (server.cuh) -- and the ids are read exactly once.
//
//   fused_direct_kernel   Localizer's key transform (:107-115), Handle::Pull (async_sgd.h:97-99,137-139,177-179),
//                         ScalarLoss::Init = SpMV::Times (learn/linear/loss.h:32-39, learn/base/spmv.h:72-91), the
//                         per-example part of Evaluate, the dual of CalcGrad (:106-115,141-156) and SpMV::TransTimes
//                         (spmv.h:95-119) into fixed-point accumulators;
//   push_sweep_kernel     Handle::Push (async_sgd.h:85-95,123-135,160-175) for every key that occurred.
//
// One block per SM; inside it, groups of 256 threads each work through their own tiles of <= 256 rows, so that one
// group's id stream overlaps another group's gathers and atomics, while all share the SM's copy of the key table.
// Per tile:  P1  every thread takes entries (coalesced 8-byte loads of the ids), transforms them, looks them up in
//                the table and leaves a 4-byte code per entry in shared memory; keys outside the table set their bit
//                in the lane's key bitmap (the push kernel's work list);
//            P2  a thread per row walks its codes: weights from shared memory (table keys) or the shard (everything
//                else), summed sequentially in CSR order -- bit-exact with the reference -- then loss, dual, AUC bin;
//            P3  the same thread adds its dual into the accumulators of the row's keys.
// What decides the design is the cost of an atomic add, measured on B200 (scripts/micro/red_bench.cu): a 64-bit RED to
// global memory takes 1.5 - 4.7 cycles of an SM per lane (60 - 180 G/s for the whole chip), 11 when a thousand
// addresses are popular and 78 when thirty-two are -- and feature ids follow a power law; a shared-memory atomic takes
// 0.2, popular address or not. So the 4 096 most frequent keys of the stream (76 % of all entries at the Criteo shape)
// are summed per block in shared memory -- a 64-bit add made of two native 32-bit atomics -- and each block stores its
// sums once, without atomics, into its own row of a [blocks x slots] array that the push kernel adds up. Only keys
// outside the table pay a RED, into the shard entry that the forward gather has just brought into L2.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <numeric>
#include <vector>

#include "direct.cuh"
#include "handle.cuh"

namespace wh {

constexpr int kDxGroup = 256;                       // threads per group = rows per tile
constexpr unsigned kDxFlag = 0x80000000u;           // code bit: a table key; the low bits hold its slot
#ifndef WH_DX_ILP
#define WH_DX_ILP 4
#endif
#ifndef WH_DX_GATHER
#define WH_DX_GATHER 12
#endif
#ifndef WH_DX_MAXNREG
#define WH_DX_MAXNREG 64      // one 1024-thread block per SM (shared memory decides that): 64 registers are free
#endif
constexpr int kDxIlp = WH_DX_ILP;                   // ids in flight per thread in P1
constexpr int kDxGather = WH_DX_GATHER;             // weight gathers in flight per thread in P2

__device__ __forceinline__ void group_bar(int g) {
  asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(kDxGroup) : "memory");
}
__device__ __forceinline__ unsigned dx_bucket(unsigned k32) { return (k32 * 0x9E3779B1u) >> (32 - 10); }
// slot of key k32 in the (shared-memory) table, or -1: a 4-bit match mask and its lowest set bit
__device__ __forceinline__ int dx_probe(const uint4* s_tags, unsigned k32) {
  const unsigned b = dx_bucket(k32);
  const uint4 t = s_tags[b];
  // at most one way matches: four compares that yield all-ones / zero, folded with logic ops (selects and a
  // find-first-set compiled to 34 instructions per probe)
  unsigned m0, m1, m2, m3;
  asm("set.eq.u32.u32 %0, %1, %2;" : "=r"(m0) : "r"(t.x), "r"(k32));
  asm("set.eq.u32.u32 %0, %1, %2;" : "=r"(m1) : "r"(t.y), "r"(k32));
  asm("set.eq.u32.u32 %0, %1, %2;" : "=r"(m2) : "r"(t.z), "r"(k32));
  asm("set.eq.u32.u32 %0, %1, %2;" : "=r"(m3) : "r"(t.w), "r"(k32));
  const unsigned way = (m1 & 1u) | (m2 & 2u) | (m3 & 3u);
  const unsigned hit = m0 | m1 | m2 | m3;                     // all ones or zero
  return (int)(((b * kDxWays) | way) | ~hit);                 // slot, or -1
}
// A 64-bit sum in shared memory out of three 32-bit ones that never carry into each other: q = w0 + 2^D w1 + 2^2D w2
// with w0, w1 the low two D-bit digits (unsigned) and w2 the arithmetic rest. D is chosen per block from the number of
// entries E in its band of rows, 2^(32-D) > E, so that no sum of w0 or w1 can reach 2^32; with |q| < 2^39 and D >= 14
// the third stays far from wrapping too (bands of up to 2^18 entries; the launch sizes its grid for 40 000). Three
// atomics that return nothing: no thread waits for one. (A 64-bit atomicAdd on shared memory compiles to a
// compare-and-swap loop; two 32-bit ones with the carry taken from the first one's returned value left every thread
// waiting on that value -- 17 % of the kernel's stall samples.)
struct DxTerm { unsigned w0, w1; int w2; };
__device__ __forceinline__ DxTerm dx_term(long long q, unsigned D) {
  DxTerm t;
  const unsigned long long u = (unsigned long long)q;
  const unsigned mask = (1u << D) - 1u;
  t.w0 = (unsigned)u & mask; t.w1 = (unsigned)(u >> D) & mask; t.w2 = (int)(q >> (2 * D));
  return t;
}
__device__ __forceinline__ void dx_smem_add(unsigned* s_acc, unsigned slot, const DxTerm& t) {
  atomicAdd(s_acc + slot, t.w0);
  atomicAdd(s_acc + kDxSlots + slot, t.w1);
  if (t.w2) atomicAdd(s_acc + 2 * kDxSlots + slot, (unsigned)t.w2);
}
__device__ __forceinline__ long long dx_smem_sum(const unsigned* s_acc, unsigned slot, unsigned D) {
  return (long long)s_acc[slot] + ((long long)s_acc[kDxSlots + slot] << D) + ((long long)(int)s_acc[2 * kDxSlots + slot] << (2 * D));
}

struct DxArgs {
  int loss;
  unsigned rows, rows_per_block, ecap;
  const size_t* offset; const uint64_t* index; const float* value; const float* label;
  KeyXform xf;
  ShardEntry* shard;        // a key's weight is shard[key].w, its accumulator shard[key].acc
  const uint4* tags; const float* hot_w;
  const uint32_t* codes;    // non-NULL: one code per entry, staged by dx_stage_kernel
  unsigned* bitmap; long long* part; unsigned* part_touched;
  float* xw; double* part_objv; unsigned* part_correct; unsigned long long* auc_hist;
  FixedScale* fs;
  long long* counters;      // the server's: [1] counts what wh_worker_wait reports as an error
  float scale; double inv;
  int phases;               // 3 = the whole step. 1 / 2: stop after P1 / P2 (timing experiments only: results are wrong)
  long long* timeline;      // WH_DX_TIMELINE: clock64 at the phase boundaries of every group's first tile, 8 per group
};
#define DX_MARK(i) do { if (a.timeline && tg == 0 && first_tile) a.timeline[((size_t)blockIdx.x * NG + g) * 8 + (i)] = clock64(); } while (0)

// P1 for entries [0, E) of a tile that starts at entry `base`: codes (and values) into shared memory. The ids of the
// next trip are in flight while this trip's are looked up.
template <bool TRAIN, bool POW2, bool HAS_VAL>
__device__ __forceinline__ void dx_stage(const DxArgs& a, const uint4* s_tags, unsigned char* s_touch, uint32_t* s_code, float* s_val,
                                         size_t base, unsigned E, unsigned tg) {
  if (a.codes) {
    // the codes were made ahead of time by dx_stage_kernel (minibatches in flight: that part of the step does not
    // depend on the model, so it runs while the previous minibatch is still in its forward / backward / push)
    const uint64_t pol = l2_policy_evict_first();
    for (unsigned e = tg; e < E; e += kDxGroup) {
      s_code[e] = ld_stream_evict_first_u32(a.codes + base + e, pol);
      if (HAS_VAL) s_val[e] = __ldcs(a.value + base + e);
    }
    return;
  }
  const unsigned mask = (unsigned)(a.xf.max_key - 1);
  const uint64_t* ids = a.index + base;
  uint64_t nxt[kDxIlp];
  const uint64_t once = l2_policy_evict_first();      // (one createpolicy per call, not one per load)
#pragma unroll
  for (int u = 0; u < kDxIlp; ++u) { const unsigned e = tg + u * kDxGroup; nxt[u] = e < E ? ld_stream_evict_first_u64(ids + e, once) : 0; }
  for (unsigned e0 = tg; e0 < E; e0 += kDxGroup * kDxIlp) {
    uint64_t id[kDxIlp];
    float v[kDxIlp];
#pragma unroll
    for (int u = 0; u < kDxIlp; ++u) {
      id[u] = nxt[u];
      const unsigned en = e0 + (kDxIlp + u) * kDxGroup;
      nxt[u] = en < E ? ld_stream_evict_first_u64(ids + en, once) : 0;
      if (HAS_VAL) { const unsigned e = e0 + u * kDxGroup; v[u] = e < E ? __ldcs(a.value + base + e) : 0.f; }
    }
#pragma unroll
    for (int u = 0; u < kDxIlp; ++u) {
      const unsigned e = e0 + u * kDxGroup;
      if (e >= E) continue;
      const unsigned k = POW2 ? ((unsigned)id[u] & mask) : (unsigned)a.xf(id[u]);   // direct addressing: keys < 2^28
      const int slot = dx_probe(s_tags, k);
      unsigned code;
      if (slot >= 0) {
        code = kDxFlag | (unsigned)slot;
        if (TRAIN) s_touch[slot] = 1;   // (same-value byte stores: no atomic needed)
      } else {
        code = k;
        if (TRAIN) atomicOr(a.bitmap + (k >> 5), 1u << (k & 31));
      }
      s_code[e] = code;
      if (HAS_VAL) s_val[e] = v[u];
    }
  }
}

// weight of a code: table keys from shared memory, the rest from the shard -- one generic load from a selected address
// (two predicated loads, one of which issues: a generic load from a selected address cost 51 instructions per entry in
//  address-space conversion and 64-bit selects, a branch around two loads 21)
// The entry of a key outside the table is wanted three times within the minibatch -- this gather, the accumulating RED,
// the push -- and the ids, codes and partial sums streaming through L2 in between would evict it: evict-last policy.
__device__ __forceinline__ float dx_weight(const ShardEntry* __restrict__ shard, unsigned s_hw32, uint32_t c, uint64_t keep) {
  float w;
  const unsigned saddr = s_hw32 + ((c & (kDxSlots - 1)) << 2);
  const ShardEntry* g = shard + (c & 0x0fffffffu);
  asm volatile("{\n\t.reg .pred p;\n\tsetp.lt.s32 p, %3, 0;\n\t@p ld.shared.f32 %0, [%1];\n\t@!p ld.global.nc.L1::no_allocate.L2::cache_hint.f32 %0, [%2], %4;\n\t}"
               : "=f"(w) : "r"(saddr), "l"(g), "r"(c), "l"(keep));
  return w;
}
// one entry's term into the accumulators (P3): the block's shared-memory sums, or the key's entry in the shard
__device__ __forceinline__ void dx_add(ShardEntry* shard, unsigned* s_acc, uint32_t c, long long q, const DxTerm& t) {
  if (c & kDxFlag) dx_smem_add(s_acc, c & (kDxSlots - 1), t);
  else atomicAdd(reinterpret_cast<unsigned long long*>(&shard[c].acc), (unsigned long long)q);
}

// A single row with more entries than a tile holds: chunk by chunk, one thread keeps the reference's summation
// order. Correctness path only (a row of > 8 000 features).
template <bool TRAIN, bool POW2, bool HAS_VAL>
__device__ void dx_long_row(const DxArgs& a, const uint4* s_tags, unsigned s_hw, unsigned char* s_touch, unsigned* s_acc,
                            uint32_t* s_code, float* s_val, float* s_bcast, unsigned row, int g, unsigned tg, float scale,
                            unsigned D, double* objv, unsigned* correct) {
  const size_t b = a.offset[row], e = a.offset[row + 1];
  float sum = 0.f;
  for (size_t c0 = b; c0 < e; c0 += a.ecap) {
    const unsigned E = (unsigned)min((size_t)a.ecap, e - c0);
    dx_stage<false, POW2, HAS_VAL>(a, s_tags, s_touch, s_code, s_val, c0, E, tg);
    group_bar(g);
    if (tg == 0)
      for (unsigned j = 0; j < E; ++j) { float w = dx_weight(a.shard, s_hw, s_code[j], l2_policy_evict_last()); if (HAS_VAL) w *= s_val[j]; sum += w; }
    group_bar(g);
  }
  float dl = 0.f;
  if (tg == 0) {
    a.xw[row] = sum;
    double o; unsigned cr;
    loss_row(a.loss, a.label[row], sum, &dl, &o, &cr);
    *objv += o; *correct += cr;
    auc_hist_add(a.auc_hist, a.label[row], sum);
    *s_bcast = dl;
  }
  group_bar(g);
  dl = *s_bcast;
  if (TRAIN) {
    for (size_t c0 = b; c0 < e; c0 += a.ecap) {
      const unsigned E = (unsigned)min((size_t)a.ecap, e - c0);
      dx_stage<true, POW2, HAS_VAL>(a, s_tags, s_touch, s_code, s_val, c0, E, tg);
      group_bar(g);
      for (unsigned j = tg; j < E; j += kDxGroup) {
        const long long q = __float2ll_rn((HAS_VAL ? dl * s_val[j] : dl) * scale);
        dx_add(a.shard, s_acc, s_code[j], q, dx_term(q, D));
      }
      group_bar(g);
    }
  }
}

// P1 as a launch of its own, for workers that keep several minibatches in flight: ids -> codes in global memory (4 bytes
// per entry), key bitmap and table-key flags as in the fused kernel. Small blocks with little shared memory, so that
// they fit beside the step kernel of the minibatch before.
template <bool TRAIN, bool POW2>
__global__ void __launch_bounds__(kDxGroup) dx_stage_kernel(const DxArgs a, size_t nnz, uint32_t* codes_out, unsigned* touched) {
  __shared__ uint4 s_tags[kDxBuckets];
  __shared__ unsigned char s_touch[kDxSlots];
  for (int i = threadIdx.x; i < kDxBuckets; i += blockDim.x) s_tags[i] = __ldg(a.tags + i);
  for (int i = threadIdx.x; i < kDxSlots / 4; i += blockDim.x) reinterpret_cast<unsigned*>(s_touch)[i] = 0u;
  __syncthreads();
  const size_t chunk = ((nnz + gridDim.x - 1) / gridDim.x + 3) & ~(size_t)3;
  const size_t lo = min(nnz, (size_t)blockIdx.x * chunk), hi = min(nnz, lo + chunk);
  if (hi > lo) dx_stage<TRAIN, POW2, false>(a, s_tags, s_touch, codes_out + lo, nullptr, lo, (unsigned)(hi - lo), threadIdx.x);
  if (TRAIN) {
    __syncthreads();
    for (int i = threadIdx.x; i < kDxSlots; i += blockDim.x) if (s_touch[i]) touched[i] = 1u;
  }
}

template <bool TRAIN, bool POW2, bool HAS_VAL, int NG>
__global__ void __maxnreg__(WH_DX_MAXNREG) fused_direct_kernel(const DxArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  // (with the codes staged ahead of time the table's tags and flags are not needed here, and their 20 KB are left
  //  to the next minibatch's dx_stage_kernel blocks, which run beside this kernel)
  const bool staged = a.codes != nullptr;
  unsigned* s_acc = reinterpret_cast<unsigned*>(smem);                     // 48 KB: the block's sums for the table keys (3 words each)
  float* s_hwp = reinterpret_cast<float*>(s_acc + 3 * kDxSlots);           // 16 KB: the table keys' weights
  uint4* s_tags = reinterpret_cast<uint4*>(s_hwp + kDxSlots);              // 16 KB
  unsigned char* s_touch = reinterpret_cast<unsigned char*>(s_tags + (staged ? 0 : kDxBuckets));   // 4 KB: table keys this block saw
  unsigned* s_off_all = reinterpret_cast<unsigned*>(s_touch + (staged ? 0 : kDxSlots));
  const unsigned s_hw = smem_u32(s_hwp);
  float* s_bcast_all = reinterpret_cast<float*>(s_off_all + NG * (kDxGroup + 4));
  uint32_t* s_code_all = reinterpret_cast<uint32_t*>(s_bcast_all + 4);
  __shared__ double s_o[32];
  __shared__ unsigned s_c[32];

  const int g = threadIdx.x / kDxGroup;
  const unsigned tg = threadIdx.x % kDxGroup, lane = threadIdx.x & 31;
  unsigned* s_off = s_off_all + g * (kDxGroup + 4);
  uint32_t* s_code = s_code_all + (size_t)g * a.ecap * (HAS_VAL ? 2 : 1);
  float* s_val = reinterpret_cast<float*>(s_code + a.ecap);

  // the push reads the fixed-point scale from device memory; for binary features the host knows it (|dual| <= 1)
  float scale = a.scale;
  if (HAS_VAL) scale = a.fs->scale;
  else if (blockIdx.x == 0 && threadIdx.x == 0) { a.fs->scale = a.scale; a.fs->inv = a.inv; }

  bool first_tile = true;
  DX_MARK(0);
  for (int i = threadIdx.x; i < kDxBuckets; i += blockDim.x) {
    if (!a.codes) s_tags[i] = __ldg(a.tags + i);
    reinterpret_cast<float4*>(s_hwp)[i] = __ldg(reinterpret_cast<const float4*>(a.hot_w) + i);
  }
  if (TRAIN) {
    for (int i = threadIdx.x; i < 3 * kDxSlots / 4; i += blockDim.x) reinterpret_cast<int4*>(s_acc)[i] = make_int4(0, 0, 0, 0);
    if (!staged) for (int i = threadIdx.x; i < kDxSlots / 4; i += blockDim.x) reinterpret_cast<unsigned*>(s_touch)[i] = 0u;
  }
  __syncthreads();

  // this block's band of rows, and this group's share of it
  const unsigned blo = min(a.rows, blockIdx.x * a.rows_per_block), bhi = min(a.rows, blo + a.rows_per_block);
  const unsigned per = (bhi - blo + NG - 1) / NG;
  unsigned row = min(bhi, blo + g * per);
  const unsigned ghi = min(bhi, row + per);

  // digit width of the shared-memory sums (dx_term): from the entries of the block's band
  const unsigned long long band_entries = a.offset[bhi] - a.offset[blo];
  const unsigned D = min(16u, (unsigned)__clzll(band_entries | 1ull) - 32u);
  if (TRAIN && (band_entries >> 18) && threadIdx.x == 0) atomicAdd(reinterpret_cast<unsigned long long*>(a.counters + 1), 1ull);   // (reported as an error)

  double objv_t = 0;
  unsigned correct_t = 0;

  DX_MARK(1);
  while (row < ghi) {   // uniform within the group
    const unsigned want = min((unsigned)kDxGroup, ghi - row);
    const size_t base = a.offset[row];
    if (tg < want) s_off[tg + 1] = (unsigned)min(a.offset[row + tg + 1] - base, (size_t)0xffffffffu);
    if (tg == 0) s_off[0] = 0;
    group_bar(g);
    unsigned lo = 0, hi = want;   // R = the most rows whose entries fit the tile
    while (lo < hi) { const unsigned mid = (lo + hi + 1) >> 1; if (s_off[mid] <= a.ecap) lo = mid; else hi = mid - 1; }
    const unsigned R = lo;
    if (R == 0) {
      dx_long_row<TRAIN, POW2, HAS_VAL>(a, s_tags, s_hw, s_touch, s_acc, s_code, s_val, s_bcast_all + g, row, g, tg, scale,
                                        D, &objv_t, &correct_t);
      row += 1;
      group_bar(g);
      continue;
    }
    const unsigned E = s_off[R];
    dx_stage<TRAIN, POW2, HAS_VAL>(a, s_tags, s_touch, s_code, s_val, base, E, tg);
    group_bar(g);
    DX_MARK(2);
    if (a.phases < 2) { row += R; continue; }

    // P2: a thread per row
    const bool live = tg < R;
    const unsigned lb = live ? s_off[tg] : 0u, len = live ? s_off[tg + 1] - s_off[tg] : 0u;
    float dl = 0.f;
    const uint64_t keep = l2_policy_evict_last();
    unsigned hbin = 0xffffffffu;   // AUC bin << 1 | positive
    if (live) {
      float sum = 0.f;
      unsigned j = 0;
      for (; j + kDxGather <= len; j += kDxGather) {   // gathers in flight, the adds stay in CSR order
        uint32_t c[kDxGather];
        float wv[kDxGather];
#pragma unroll
        for (int u = 0; u < kDxGather; ++u) c[u] = s_code[lb + j + u];
#pragma unroll
        for (int u = 0; u < kDxGather; ++u) wv[u] = dx_weight(a.shard, s_hw, c[u], keep);
#pragma unroll
        for (int u = 0; u < kDxGather; ++u) sum += HAS_VAL ? wv[u] * s_val[lb + j + u] : wv[u];
      }
      for (; j < len; ++j) {
        float wv = dx_weight(a.shard, s_hw, s_code[lb + j], keep);
        if (HAS_VAL) wv *= s_val[lb + j];
        sum += wv;
      }
      const unsigned r = row + tg;
      const float lab = a.label[r];
      a.xw[r] = sum;   // rows without entries: Xw_ starts zeroed in ScalarLoss::Init
      double o; unsigned cr;
      loss_row(a.loss, lab, sum, &dl, &o, &cr);
      objv_t += o; correct_t += cr;
      hbin = ((ordered_bits(sum) >> 16) << 1) | (lab > 0 ? 1u : 0u);
    }
    {
      // AUC histogram: rows of a warp that fall into the same bin (an untrained or sparse model scores most rows
      // alike) add with one RED
      const unsigned peers = __match_any_sync(0xffffffffu, hbin);
      if (hbin != 0xffffffffu && (int)lane == __ffs(peers) - 1)
        atomicAdd(a.auc_hist + (hbin >> 1), ((hbin & 1u) ? (1ull << 32) : 1ull) * (unsigned long long)__popc(peers));
    }
    DX_MARK(3);
    if (TRAIN && a.phases >= 3) {
      // P3: a thread per row again
      if (live) {
        long long q = __float2ll_rn(dl * scale);
        DxTerm t = dx_term(q, D);   // binary features: the row's term is the same for all its entries
        for (unsigned j = 0; j < len; ++j) {
          if (HAS_VAL) { q = __float2ll_rn(dl * s_val[lb + j] * scale); t = dx_term(q, D); }
          dx_add(a.shard, s_acc, s_code[lb + j], q, t);
        }
      }
    }
    DX_MARK(5);
    group_bar(g);   // the tile's shared memory is free
    DX_MARK(6);
    first_tile = false;
    row += R;
  }

  // the block's evaluation partial sums
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    objv_t += __shfl_down_sync(0xffffffffu, objv_t, o);
    correct_t += __shfl_down_sync(0xffffffffu, correct_t, o);
  }
  if (lane == 0) { s_o[threadIdx.x >> 5] = objv_t; s_c[threadIdx.x >> 5] = correct_t; }
  __syncthreads();   // also: every group is done with its tiles, the shared-memory sums are final
  if (threadIdx.x == 0) {
    double o = 0; unsigned cc = 0;
    for (int k = 0; k < NG * kDxGroup / 32; ++k) { o += s_o[k]; cc += s_c[k]; }
    a.part_objv[blockIdx.x] = o; a.part_correct[blockIdx.x] = cc;
  }
  if (TRAIN) {
    // this block's row of the partial sums: plain, coalesced stores -- the push kernel adds the rows up
    long long* dst = a.part + (size_t)blockIdx.x * kDxSlots;
    for (int i = threadIdx.x; i < kDxSlots; i += blockDim.x) {
      dst[i] = dx_smem_sum(s_acc, i, D);
      const unsigned seen = __ballot_sync(0xffffffffu, !staged && s_touch[i] != 0);   // (staged: dx_stage_kernel kept the flags)
      if (lane == 0) a.part_touched[(size_t)blockIdx.x * (kDxSlots / 32) + (i >> 5)] = seen;
    }
  }
  first_tile = true;
  DX_MARK(7);
}

// ZPush. Blocks [0, hot_blocks) take the table's slots, the rest sweep the key bitmap; every key found gets one
// Handle::Push with its accumulated gradient, and accumulator and bit are left zero for the next minibatch.
// A key outside the table costs one 32-byte entry {w, z, n | accumulator}, already in L2: the forward gather and the
// accumulating REDs of the kernel before brought it in.
__device__ __forceinline__ int push_entry(const HandleParams& h, ShardEntry* e, double inv, float mul) {
  int4* p = reinterpret_cast<int4*>(e);
  const int4 lo = __ldcg(p), hi = __ldcg(p + 1);
  const float old_w = __int_as_float(lo.x);
  float z = __int_as_float(lo.y), n = __int_as_float(lo.z);
  const long long s = ((long long)hi.y << 32) | (unsigned)hi.x;
  const float nw = handle_update(h, old_w, &z, &n, (float)((double)s * inv) * mul);
  p[0] = make_int4(__float_as_int(nw), __float_as_int(z), __float_as_int(n), 0);
  p[1] = make_int4(0, 0, 0, 0);
  return nnz_delta(old_w, nw);
}
__global__ void __launch_bounds__(256)
push_sweep_kernel(HandleParams h, ShardView t, unsigned* __restrict__ bitmap, unsigned words,
                  const unsigned* __restrict__ tags, const long long* __restrict__ part, const unsigned* __restrict__ part_touched,
                  unsigned* __restrict__ staged_touched, unsigned nparts, float* __restrict__ hot_w, const FixedScale* __restrict__ fs, float mul,
                  long long* counters, unsigned hot_blocks) {
  const double inv = fs->inv;
  int delta = 0;
  if (blockIdx.x < hot_blocks) {
    // 32 slots per block, 8 threads per slot: each adds up an eighth of the blocks' partial sums (coalesced across the
    // slots), then the eight meet through shared memory. (A single thread per slot walked 148 rows one L2 latency
    // after the other: 19 us.)
    __shared__ long long s_sum[8][32];
    __shared__ unsigned s_seen[8][32];
    const unsigned sl = threadIdx.x & 31, ch = threadIdx.x >> 5, slot = blockIdx.x * 32 + sl;
    long long s = 0;
    unsigned seen = 0;
    for (unsigned b = ch; b < nparts; b += 8) {
      s += __ldcg(part + (size_t)b * kDxSlots + slot);
      seen |= __ldcg(part_touched + (size_t)b * (kDxSlots / 32) + (slot >> 5));
    }
    s_sum[ch][sl] = s; s_seen[ch][sl] = seen;
    __syncthreads();
    if (ch == 0) {
#pragma unroll
      for (int c = 1; c < 8; ++c) { s += s_sum[c][sl]; seen |= s_seen[c][sl]; }
      bool touched = (seen >> (slot & 31)) & 1u;
      if (staged_touched && staged_touched[slot]) { touched = true; staged_touched[slot] = 0u; }   // (flags of dx_stage_kernel)
      if (touched) {
        const unsigned k = tags[slot];
        float nw;
        delta = handle_push(h, t, k, (float)((double)s * inv) * mul, &nw);
        hot_w[slot] = nw;
      }
    }
  } else {
    // a bitmap word per thread. First every key of the word sends for its entry (prefetch: nothing waits), then the
    // keys are updated one after the other -- walking them load by load cost a DRAM latency per key.
    const unsigned nthreads = (gridDim.x - hot_blocks) * blockDim.x;
    for (unsigned i = (blockIdx.x - hot_blocks) * blockDim.x + threadIdx.x; i < words; i += nthreads) {
      const unsigned word = __ldcs(bitmap + i);
      if (!word) continue;
      bitmap[i] = 0u;
      const unsigned kb = i << 5;
      for (unsigned bits = word; bits; bits &= bits - 1)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(t.e + kb + (__ffs(bits) - 1)));
      for (unsigned bits = word; bits; bits &= bits - 1)
        delta += push_entry(h, t.e + kb + (__ffs(bits) - 1), inv, mul);
    }
  }
  commit_counters(counters, delta, 0);
}

__global__ void direct_refresh_w_kernel(const unsigned* __restrict__ tags, const ShardEntry* __restrict__ shard, float* hot_w) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= kDxSlots) return;
  const unsigned k = tags[p];
  hot_w[p] = k == kDxEmpty ? 0.f : __ldg(&shard[k].w);
}

int direct_table_refresh_w(whContext* ctx, DirectTable* t, const ShardEntry* shard) {
  WH_LAUNCH(direct_refresh_w_kernel, kDxSlots / 256, 256, 0, ctx->stream, t->tags.as<unsigned>(), shard, t->hot_w.as<float>());
  return WH_OK;
}

// Most frequent keys first, each into its bucket's first free way.
int direct_table_build(whContext* ctx, DirectTable* t, const uint64_t* keys, const uint32_t* counts, size_t n) {
  std::vector<uint32_t> image(kDxSlots, kDxEmpty), order(n);
  std::iota(order.begin(), order.end(), 0u);
  const size_t want = std::min<size_t>(n, (size_t)kDxSlots * 2);   // candidates: twice the slots, so that the table fills up
  std::partial_sort(order.begin(), order.begin() + want, order.end(), [&](uint32_t x, uint32_t y) {
    return counts[x] != counts[y] ? counts[x] > counts[y] : keys[x] < keys[y];
  });
  int placed = 0;
  unsigned long long covered = 0, total = 0;
  for (size_t i = 0; i < n; ++i) total += counts[i];
  const char* off = getenv("WH_DX_TABLE");   // WH_DX_TABLE=0: an empty table, every key takes the general route (bench.py's comparison)
  for (size_t i = 0; i < want && !(off && atoi(off) == 0); ++i) {
    const uint64_t k = keys[order[i]];
    if (counts[order[i]] < 2 || k >= (1ull << 28)) continue;
    const unsigned k32 = (unsigned)k, b = (k32 * 0x9E3779B1u) >> (32 - 10);
    for (int w = 0; w < kDxWays; ++w) {
      const int slot = (int)b * kDxWays + w;
      if (image[slot] != kDxEmpty) continue;
      image[slot] = k32;
      ++placed;
      covered += counts[order[i]];
      break;
    }
  }
  t->placed = placed;   // (an empty table is a valid table: every key then takes the general route)
  t->coverage = total ? (double)covered / (double)total : 0.0;
  WH_TRY(t->tags.reserve((size_t)kDxSlots * 4)); WH_TRY(t->hot_w.reserve((size_t)kDxSlots * 4));
  WH_CUDA(cudaMemcpyAsync(t->tags.p, image.data(), (size_t)kDxSlots * 4, cudaMemcpyHostToDevice, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));   // the host vector goes out of scope
  t->ready = true;
  return WH_OK;
}

int direct_lane_prepare(whContext* ctx, DirectLane* l, uint64_t max_key) {
  if (l->keys >= max_key && l->bitmap.p) return WH_OK;
  const size_t words = align_up(div_up(max_key, 32), 4);
  WH_TRY(l->bitmap.reserve(words * 4));
  WH_TRY(l->part.reserve((size_t)kDxMaxBlocks * kDxSlots * 8)); WH_TRY(l->part_touched.reserve((size_t)kDxMaxBlocks * (kDxSlots / 32) * 4));
  WH_CUDA(cudaMemsetAsync(l->bitmap.p, 0, l->bitmap.cap, ctx->stream));
  l->keys = max_key;
  return WH_OK;
}

template <bool TRAIN, bool POW2, bool HAS_VAL, int NG>
static int dx_launch2(unsigned grid, size_t smem, cudaStream_t stream, const DxArgs& a) {
  WH_CUDA(cudaFuncSetAttribute(fused_direct_kernel<TRAIN, POW2, HAS_VAL, NG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  WH_LAUNCH((fused_direct_kernel<TRAIN, POW2, HAS_VAL, NG>), grid, NG * kDxGroup, smem, stream, a);
  return WH_OK;
}
template <bool TRAIN, bool POW2, bool HAS_VAL>
static int dx_launch(int ng, unsigned grid, size_t smem, cudaStream_t stream, const DxArgs& a) {
  if (ng == 4) return dx_launch2<TRAIN, POW2, HAS_VAL, 4>(grid, smem, stream, a);
  return dx_launch2<TRAIN, POW2, HAS_VAL, 3>(grid, smem, stream, a);
}
static int dx_dispatch(bool train, bool pow2, bool has_val, int ng, unsigned grid, size_t smem, cudaStream_t stream, const DxArgs& a) {
#define WH_DX(T, P, V) return dx_launch<T, P, V>(ng, grid, smem, stream, a)
  if (train) {
    if (pow2) { if (has_val) WH_DX(true, true, true); else WH_DX(true, true, false); }
    else { if (has_val) WH_DX(true, false, true); else WH_DX(true, false, false); }
  } else {
    if (pow2) { if (has_val) WH_DX(false, true, true); else WH_DX(false, true, false); }
    else { if (has_val) WH_DX(false, false, true); else WH_DX(false, false, false); }
  }
#undef WH_DX
}

int direct_fused_launch(whContext* ctx, const DirectStep& st, const DirectTable& t, DirectLane& l, unsigned* nparts) {
  static int ng_env = -1;
  if (ng_env < 0) { const char* e = getenv("WH_DX_GROUPS"); ng_env = e ? atoi(e) : 4; if (ng_env != 3) ng_env = 4; }
  int ecap_env = 0;
  { const char* e = getenv("WH_DX_ECAP"); ecap_env = e ? atoi(e) : 0; }   // tests: small tiles, rows longer than a tile
  const int ng = ng_env;
  // one block per SM; more (a further wave) only where a block would otherwise add 65 536 entries or more, which the
  // shared-memory sums could not hold (dx_smem_add)
  size_t waves = 1;
  while (st.nnz / ((size_t)ctx->num_sms * waves) * 3 / 2 >= 60000 && (size_t)ctx->num_sms * (waves + 1) <= (size_t)kDxMaxBlocks) ++waves;
  const unsigned grid = (unsigned)std::min<size_t>(std::min<size_t>((size_t)ctx->num_sms * waves, kDxMaxBlocks), div_up(st.rows, 32));
  const unsigned rpb = (unsigned)div_up(st.rows, grid);
  const bool has_val = st.value != nullptr;
  // shared memory: table (tags 16 KB, sums 32 KB, weights 16 KB) | offsets | codes (+ values). A group's tile holds up
  // to 256 rows; size the code area for the average row (longer tiles are cut short, see the R search in the kernel).
  const size_t fixed = (size_t)kDxSlots * (12 + 4 + (st.codes ? 0 : 4 + 1)) + (size_t)ng * (kDxGroup + 4) * 4 + 16;
  const size_t budget = 227 * 1024 - 1024;   // static shared memory of the kernel comes on top
  const size_t per_entry = has_val ? 8 : 4;
  const size_t avg_len = std::max<size_t>(1, div_up(st.nnz, st.rows));
  const size_t tile_rows = std::min<size_t>(kDxGroup, div_up(rpb, ng));
  size_t ecap = std::min<size_t>((budget - fixed) / ng / per_entry, tile_rows * avg_len * (st.codes ? 9 : 10) / 8 + 64);
  if (ecap_env > 0) ecap = (size_t)ecap_env;
  ecap = std::max<size_t>(64, ecap & ~(size_t)3);
  WH_REQUIRE(fixed + ecap * per_entry * ng <= budget, "direct_fused_launch: %zu bytes of shared memory", fixed + ecap * per_entry * ng);
  const size_t smem = fixed + ecap * per_entry * ng;
  DxArgs a;
  a.loss = st.loss; a.rows = (unsigned)st.rows; a.rows_per_block = rpb; a.ecap = (unsigned)ecap;
  a.offset = st.offset; a.index = st.index; a.value = st.value; a.label = st.label; a.xf = st.xf; a.shard = st.shard;
  a.tags = t.tags.as<uint4>(); a.hot_w = t.hot_w.as<float>(); a.codes = st.codes;
  a.bitmap = l.bitmap.as<unsigned>(); a.part = l.part.as<long long>(); a.part_touched = l.part_touched.as<unsigned>();
  a.xw = st.xw; a.part_objv = st.part_objv; a.part_correct = st.part_correct; a.auc_hist = st.auc_hist; a.fs = st.fs; a.counters = st.counters;
  unit_terms_scale(st.nnz, &a.scale, &a.inv);
  a.phases = 3;
  if (const char* e = getenv("WH_DX_DEBUG_PHASES")) {   // timing experiments: the step is cut short and its results are wrong
    a.phases = atoi(e);
    static bool warned = false;
    if (!warned) { fprintf(stderr, "wormhole_b200: WH_DX_DEBUG_PHASES=%d -- results are WRONG, timing experiment only\n", a.phases); warned = true; }
  }
  a.timeline = nullptr;
  static long long* tl_dev = nullptr;
  const char* tl_env = getenv("WH_DX_TIMELINE");
  if (tl_env) {
    if (!tl_dev) { WH_CUDA(cudaMalloc(&tl_dev, (size_t)1024 * 4 * 8 * 8)); }
    WH_CUDA(cudaMemsetAsync(tl_dev, 0, (size_t)1024 * 4 * 8 * 8, ctx->stream));
    a.timeline = tl_dev;
  }
  const bool pow2 = st.xf.mode == 1;
  *nparts = grid;
  if (st.train) l.blocks = grid;
  WH_TRY(dx_dispatch(st.train, pow2, has_val, ng, grid, smem, ctx->stream, a));
  if (tl_env) {   // profiling aid: where a group's time goes, in cycles (serialises the stream)
    std::vector<long long> h((size_t)grid * ng * 8);
    WH_CUDA(cudaMemcpyAsync(h.data(), tl_dev, h.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
    WH_CUDA(cudaStreamSynchronize(ctx->stream));
    double d[7] = {0}; double mx[7] = {0}; int cnt = 0;
    for (unsigned gi = 0; gi < grid * (unsigned)ng; ++gi) {
      const long long* tt = &h[(size_t)gi * 8];
      if (!tt[0] || !tt[7] || !tt[2]) continue;
      ++cnt;
      const int idx[8] = {0, 1, 2, 3, 3, 5, 6, 7};
      for (int i = 0; i < 7; ++i) { const double v = (double)(tt[idx[i + 1]] - tt[idx[i]]); d[i] += v; if (v > mx[i]) mx[i] = v; }
    }
    if (cnt) fprintf(stderr, "dx timeline (cycles, avg | max over %d groups): prologue %.0f|%.0f  P1 %.0f|%.0f  P2 %.0f|%.0f  P3 %.0f|%.0f  bar %.0f|%.0f  tail %.0f|%.0f\n",
                     cnt, d[0] / cnt, mx[0], d[1] / cnt, mx[1], d[2] / cnt, mx[2], d[4] / cnt, mx[4], d[5] / cnt, mx[5], d[6] / cnt, mx[6]);
  }
  return WH_OK;
}

int direct_stage_launch(whContext* ctx, const DirectStep& st, const DirectTable& t, DirectLane& l) {
  WH_TRY(l.codes.reserve(st.nnz * 4 + 64));
  if (!l.touched.p) {
    WH_TRY(l.touched.reserve((size_t)kDxSlots * 4));
    WH_CUDA(cudaMemsetAsync(l.touched.p, 0, (size_t)kDxSlots * 4, ctx->stream));
  }
  DxArgs a;
  memset(&a, 0, sizeof(a));
  a.index = st.index; a.xf = st.xf; a.tags = t.tags.as<uint4>(); a.bitmap = l.bitmap.as<unsigned>(); a.codes = nullptr;
  // two small blocks per SM: they have to fit into what the step kernel of the minibatch before leaves free
  const unsigned grid = (unsigned)std::min<size_t>((size_t)ctx->num_sms * 2, div_up(st.nnz, 1024));
  uint32_t* codes = l.codes.as<uint32_t>();
  unsigned* touched = l.touched.as<unsigned>();
  const bool pow2 = st.xf.mode == 1;
  if (st.train) {
    if (pow2) WH_LAUNCH((dx_stage_kernel<true, true>), grid, kDxGroup, 0, ctx->stream, a, st.nnz, codes, touched);
    else WH_LAUNCH((dx_stage_kernel<true, false>), grid, kDxGroup, 0, ctx->stream, a, st.nnz, codes, touched);
  } else {
    if (pow2) WH_LAUNCH((dx_stage_kernel<false, true>), grid, kDxGroup, 0, ctx->stream, a, st.nnz, codes, touched);
    else WH_LAUNCH((dx_stage_kernel<false, false>), grid, kDxGroup, 0, ctx->stream, a, st.nnz, codes, touched);
  }
  return WH_OK;
}

int direct_push_launch(whServer* s, cudaStream_t stream, const DirectTable& t, const DirectLane& l, const FixedScale* fs,
                       float mul, int num_sms) {
  HandleParams h = make_params(s, true);   // one push message: SGDHandle::Start
  const unsigned words = (unsigned)div_up(l.keys, 32);
  const unsigned hot_blocks = kDxSlots / 32;
  const unsigned sweep = (unsigned)std::min<size_t>(div_up(words, 256), (size_t)num_sms * 8);
  WH_LAUNCH(push_sweep_kernel, hot_blocks + sweep, 256, 0, stream, h, s->view(), l.bitmap.as<unsigned>(), words,
            t.tags.as<unsigned>(), l.part.as<long long>(), l.part_touched.as<unsigned>(), l.touched.as<unsigned>(), l.blocks,
            t.hot_w.as<float>(), fs, mul,
            s->counters.as<long long>(), hot_blocks);
  return WH_OK;
}

}  // namespace wh
