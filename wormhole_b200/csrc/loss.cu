// Piece (3): ScalarLoss / LogitLoss / SquareHingeLoss (learn/linear/loss.h) and BinClassEval
// (learn/base/binary_class_evaluation.h) on the device.
//
// Arithmetic notes (fp32 parity with the reference's glibc build):
//  * exp/log are evaluated in fp64 and rounded to fp32: that is the correctly rounded fp32
//    result in all but ~1e-9 of cases, which is also what glibc's expf/logf deliver; every other
//    step (the +1, the division, the products) is a single IEEE fp32 operation, compiled with
//    -fmad=false so that nothing is contracted.
//  * sums over examples (objective) are taken in fp64 in a fixed order; the reference sums in
//    fp32 under an OpenMP reduction whose order depends on the thread count.
//  * AUC is exact integer pair counting. Equal scores: the reference's std::sort leaves their
//    order unspecified (binary_class_evaluation.h:25-26); here tied (pos,neg) pairs count 1/2,
//    and the two extremes are reported as auc_lo / auc_hi.
#include <cooperative_groups.h>
#include <stdlib.h>

#include "loss.cuh"
#include "scan.cuh"
#include "spmv.cuh"

namespace wh {


__global__ void eval_reset_kernel(EvalOut* out) {
  out->objv = 0; out->logloss = 0; out->correct = 0; out->npos = 0; out->ngroups = 0;
  out->area_lo = 0; out->ties = 0;
}

// learn/linear/loss.h:109-113 (logit dual), :145-149 (squared hinge dual), :131-138 and
// binary_class_evaluation.h:66-74 (objectives), :40-51 (accuracy), :53-64 (logloss).
__global__ void __launch_bounds__(256)
loss_forward_kernel(int loss, size_t rows, const float* __restrict__ label, const float* __restrict__ xw,
                    float* __restrict__ dual, double* part_objv, double* part_ll, unsigned* part_correct,
                    int want_logloss) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  double objv = 0, ll = 0;
  unsigned correct = 0;
  if (i < rows) {
    const float lab = label[i], p = xw[i];
    float dl;
    loss_row(loss, lab, p, &dl, &objv, &correct);
    if (dual) dual[i] = dl;
    if (want_logloss) {
      float yy = lab > 0 ? 1.f : 0.f;
      float q = 1.f / (1.f + exp_rn(-p));
      q = q < 1e-10f ? 1e-10f : q;
      ll = (double)(yy * log_rn(q) + (1.f - yy) * log_rn(1.f - q));
    }
  }
  // fixed-order block reduction
  __shared__ double s_o[8], s_l[8];
  __shared__ unsigned s_c[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    objv += __shfl_down_sync(0xffffffffu, objv, o);
    ll += __shfl_down_sync(0xffffffffu, ll, o);
    correct += __shfl_down_sync(0xffffffffu, correct, o);
  }
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { s_o[warp] = objv; s_l[warp] = ll; s_c[warp] = correct; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double o = 0, l = 0; unsigned c = 0;
    for (int w = 0; w < 8; ++w) { o += s_o[w]; l += s_l[w]; c += s_c[w]; }
    part_objv[blockIdx.x] = o; part_ll[blockIdx.x] = l; part_correct[blockIdx.x] = c;
  }
}

__global__ void __launch_bounds__(256)
loss_reduce_kernel(unsigned nparts, const double* part_objv, const double* part_ll,
                   const unsigned* part_correct, EvalOut* out, FixedScale* fs, float fs_scale, double fs_inv) {
  __shared__ double s_o[256], s_l[256];
  __shared__ unsigned long long s_c[256];
  double o = 0, l = 0; unsigned long long c = 0;
  for (unsigned i = threadIdx.x; i < nparts; i += 256) { o += part_objv[i]; if (part_ll) l += part_ll[i]; c += part_correct[i]; }
  s_o[threadIdx.x] = o; s_l[threadIdx.x] = l; s_c[threadIdx.x] = c;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) { s_o[threadIdx.x] += s_o[threadIdx.x + s]; s_l[threadIdx.x] += s_l[threadIdx.x + s]; s_c[threadIdx.x] += s_c[threadIdx.x + s]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    out->objv = s_o[0]; out->logloss = s_l[0]; out->correct = s_c[0];
    out->npos = 0; out->ngroups = 0; out->area_lo = 0; out->ties = 0;   // the AUC kernel adds into these
    if (fs) { fs->scale = fs_scale; fs->inv = fs_inv; }
  }
}

int loss_reduce_launch(whContext* ctx, unsigned nparts, const double* po, const double* pl, const unsigned* pc,
                       EvalOut* out, FixedScale* fs, size_t nnz) {
  float sc = 1.f; double inv = 1.0;
  if (fs) unit_terms_scale(nnz, &sc, &inv);
  WH_LAUNCH(loss_reduce_kernel, 1, 256, 0, ctx->stream, nparts, po, pl, pc, out, fs, sc, inv);
  return WH_OK;
}

int loss_forward_launch(whContext* ctx, int loss, size_t rows, const float* label, const float* xw,
                        float* dual, EvalOut* out, int want_logloss) {
  WH_LAUNCH(eval_reset_kernel, 1, 1, 0, ctx->stream, out);
  if (rows == 0) return WH_OK;
  const unsigned nb = (unsigned)div_up(rows, 256);
  WH_TRY(ctx->stage[7].reserve((size_t)nb * 24 + 64));
  double* po = ctx->stage[7].as<double>();
  double* pl = po + nb;
  unsigned* pc = reinterpret_cast<unsigned*>(pl + nb);
  WH_LAUNCH(loss_forward_kernel, nb, 256, 0, ctx->stream, loss, rows, label, xw, dual, po, pl, pc, want_logloss);
  WH_LAUNCH(loss_reduce_kernel, 1, 256, 0, ctx->stream, nb, po, pl, pc, out, (FixedScale*)nullptr, 1.f, 1.0);
  return WH_OK;
}

// ---------------------------------------------------------------------------------- AUC --


// key = ordered(score) << 1 | (label > 0): ascending by score, negatives first among equals.
__global__ void __launch_bounds__(256)
auc_keys_kernel(size_t n, const float* __restrict__ label, const float* __restrict__ xw, uint64_t* keys,
                uint32_t* vals, SortPlan* plan) {
  unsigned long long o = 0ull, a = ~0ull;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    uint64_t k = ((uint64_t)ordered_bits(xw[i]) << 1) | (label[i] > 0 ? 1u : 0u);
    keys[i] = k; vals[i] = 0;
    o |= k; a &= k;
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) { o |= __shfl_xor_sync(0xffffffffu, o, s); a &= __shfl_xor_sync(0xffffffffu, a, s); }
  if ((threadIdx.x & 31) == 0) { atomicOr(&plan->bits_or, o); atomicAnd(&plan->bits_and, a); }
}

// scanned value: (is positive) << 32 | (starts a new score group)
struct AucFlag {
  SortBufs b;
  __device__ unsigned long long operator()(size_t i) const {
    const uint64_t* k = sorted_keys(b);
    unsigned long long head = (i == 0 || (k[i] >> 1) != (k[i - 1] >> 1)) ? 1ull : 0ull;
    return ((k[i] & 1ull) << 32) | head;
  }
};
// at each group start remember how many positives / negatives precede it
struct AucApply {
  uint32_t* gpos; uint32_t* gneg;
  __device__ void operator()(size_t i, unsigned long long v, unsigned long long prefix) const {
    if (v & 1ull) {
      unsigned g = (unsigned)(prefix & 0xffffffffull);
      unsigned pos_before = (unsigned)(prefix >> 32);
      gpos[g] = pos_before;
      gneg[g] = (unsigned)i - pos_before;
    }
  }
};
// per score group g: negatives in g each see gpos[g] positives strictly below; ties = P_g * N_g
__global__ void __launch_bounds__(256)
auc_groups_kernel(size_t n, const uint32_t* __restrict__ gpos, const uint32_t* __restrict__ gneg,
                  const unsigned long long* total, EvalOut* out) {
  const unsigned long long tot = *total;
  const unsigned ngroups = (unsigned)(tot & 0xffffffffull), P = (unsigned)(tot >> 32);
  const unsigned N = (unsigned)n - P;
  unsigned long long area = 0, ties = 0;
  for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += (size_t)gridDim.x * blockDim.x) {
    unsigned p0 = gpos[g], n0 = gneg[g];
    unsigned p1 = (g + 1 < ngroups) ? gpos[g + 1] : P, n1 = (g + 1 < ngroups) ? gneg[g + 1] : N;
    area += (unsigned long long)(n1 - n0) * p0;
    ties += (unsigned long long)(n1 - n0) * (p1 - p0);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { area += __shfl_xor_sync(0xffffffffu, area, o); ties += __shfl_xor_sync(0xffffffffu, ties, o); }
  if ((threadIdx.x & 31) == 0) { atomicAdd(&out->area_lo, area); atomicAdd(&out->ties, ties); }
  if (blockIdx.x == 0 && threadIdx.x == 0) { out->npos = P; out->ngroups = ngroups; }
}

int auc_launch_multiblock(whContext* ctx, size_t n, const float* label, const float* xw, EvalOut* out) {
  if (n == 0) return WH_OK;
  WH_REQUIRE(n < (size_t)INT32_MAX, "AUC: %zu examples exceed 32-bit counters", n);
  SortBufs b;
  WH_TRY(ctx->sort_tmp[0].reserve(n * 8)); WH_TRY(ctx->sort_tmp[1].reserve(n * 8));
  WH_TRY(ctx->sort_tmp[2].reserve(n * 4)); WH_TRY(ctx->sort_tmp[3].reserve(n * 4));
  WH_TRY(ctx->sort_tmp[4].reserve(sort_counts_bytes(n) + sizeof(SortPlan) + 256));
  WH_TRY(ctx->sort_tmp[5].reserve((n + 1) * 8));
  WH_TRY(ctx->scan_tmp.reserve(div_up(n, kScanTile) * sizeof(unsigned long long) + 256));
  b.keys[0] = ctx->sort_tmp[0].as<uint64_t>(); b.keys[1] = ctx->sort_tmp[1].as<uint64_t>();
  b.vals[0] = ctx->sort_tmp[2].as<uint32_t>(); b.vals[1] = ctx->sort_tmp[3].as<uint32_t>();
  b.plan = ctx->sort_tmp[4].as<SortPlan>();
  b.counts = reinterpret_cast<uint32_t*>(ctx->sort_tmp[4].as<char>() + 256);
  uint32_t* gpos = ctx->sort_tmp[5].as<uint32_t>();
  uint32_t* gneg = gpos + n + 1;
  unsigned long long* sums = ctx->scan_tmp.as<unsigned long long>();
  unsigned long long* total = reinterpret_cast<unsigned long long*>(ctx->small.as<char>() + 2048);
  WH_TRY(sort_plan_reset(ctx, b.plan));
  const unsigned grid = (unsigned)std::min<size_t>(div_up(n, 256), (size_t)ctx->num_sms * 8);
  WH_LAUNCH(auc_keys_kernel, grid, 256, 0, ctx->stream, n, label, xw, b.keys[0], b.vals[0], b.plan);
  WH_TRY(radix_sort_pairs(ctx, b, n));
  WH_TRY(scan_and_apply<unsigned long long>(ctx, AucFlag{b}, AucApply{gpos, gneg}, n, sums, total));
  WH_LAUNCH(auc_groups_kernel, grid, 256, 0, ctx->stream, n, gpos, gneg, total, out);
  return WH_OK;
}


// ---- AUC without a sort: two-level counting on the ordered score bits ---------------------------------------
// AUC needs, for every negative example, the number of positives with a smaller score (and with an equal one).
// A sort delivers that, but a sort of 100 000 keys is a chain of passes with barriers in between (the one-cluster
// kernel below: 165 us on 8 SMs). Counting does not need the order:
//   hist     bin = top 16 bits of the order-preserving score key (sign, exponent, 7 mantissa bits); one 64-bit RED per
//            example into {positives << 32 | negatives} of its bin. (The worker's fused kernel does this in its
//            epilogue, where the score is in a register.)
//   scan     one block walks the 65 536 bins: positives below each bin (which settles every pair whose members fall
//            into different bins: sum over bins of negatives x positives below), and the slice of `pos_low` that
//            will hold the bin's positives; the histogram is zeroed on the way for the next minibatch.
//   scatter  every positive drops the low 16 bits of its key into its bin's slice (order within the bin does not
//            matter), and the bin remembers the smallest and largest of them.
//   count    every negative compares its low 16 bits with the positives of its own bin -- 60 on average at the
//            Criteo shape, because the binning is logarithmic like the floats themselves; a bin whose positives
//            are all equal (the all-zero scores of an untrained model) is answered from min == max.
// All integer: bit-reproducible, tie bounds kept. Used up to 2^20 examples; the multi-launch sort serves beyond.
constexpr unsigned kAucBins = 1u << 16;
constexpr unsigned kAucHeavy = 64;          // bins with more positives than this are sorted, and searched instead of scanned
constexpr unsigned kAucSortMax = 1u << 16;  // ... if they fit one block's shared memory
constexpr int kAucScanThreads = 1024, kAucScanBlocks = kAucBins / kAucScanThreads;
struct AucScratch {            // one per context, zeroed when allocated
  unsigned long long* hist;    // [kAucBins]  positives << 32 | negatives; all zero between minibatches
  unsigned* begin;             // [kAucBins]  first slot of the bin's positives in pos_low
  unsigned* cursor;            // [kAucBins]  next free slot while scattering; end of the slice afterwards
  unsigned* lo_min;            // [kAucBins]  smallest / largest low half among the bin's positives
  unsigned* lo_max;
  unsigned* heavy;             // [kAucBins]  bins with more than kAucHeavy positives
  unsigned long long* cross;   // [kAucScanBlocks]  pairs settled by the bins alone, per scan block
  unsigned* nheavy;            // zero between minibatches
};
constexpr size_t kAucScratchBytes = (size_t)kAucBins * (8 + 5 * 4) + kAucScanBlocks * 8 + 64;
static AucScratch auc_scratch(void* base) {
  AucScratch a;
  a.hist = static_cast<unsigned long long*>(base);
  a.begin = reinterpret_cast<unsigned*>(a.hist + kAucBins);
  a.cursor = a.begin + kAucBins; a.lo_min = a.cursor + kAucBins; a.lo_max = a.lo_min + kAucBins; a.heavy = a.lo_max + kAucBins;
  a.cross = reinterpret_cast<unsigned long long*>(a.heavy + kAucBins);
  a.nheavy = reinterpret_cast<unsigned*>(a.cross + kAucScanBlocks);
  return a;
}

__global__ void __launch_bounds__(256)
auc_hist_kernel(size_t n, const float* __restrict__ label, const float* __restrict__ xw, unsigned long long* hist) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) auc_hist_add(hist, label[i], xw[i]);
}

// One bin per thread. A block first adds up the positives of all bins below its own 1 024 (a strided, coalesced read
// of at most 512 KB out of L2 -- cheaper than a second launch or a grid barrier), then scans its own.
__global__ void __launch_bounds__(kAucScanThreads)
auc_scan_kernel(AucScratch a, unsigned nparts, const double* part_objv, const unsigned* part_correct, EvalOut* out) {
  const unsigned t = threadIdx.x, b = blockIdx.x * kAucScanThreads + t;
  const unsigned long long v = a.hist[b];
  unsigned below = 0;
  for (unsigned i = t; i < blockIdx.x * kAucScanThreads; i += kAucScanThreads) below += (unsigned)(__ldcg(a.hist + i) >> 32);
  unsigned base;
  block_exclusive_scan<unsigned>(below, &base);
  const unsigned p = (unsigned)(v >> 32), ng = (unsigned)v;
  unsigned tot;
  const unsigned pos_below = base + block_exclusive_scan<unsigned>(p, &tot);
  a.begin[b] = pos_below; a.cursor[b] = pos_below;
  if (p) { a.lo_min[b] = 0xffffffffu; a.lo_max[b] = 0u; }
  if (p > kAucHeavy) a.heavy[atomicAdd(a.nheavy, 1u)] = b;
  unsigned long long cross = (unsigned long long)ng * pos_below;
  __shared__ unsigned long long s_x[kAucScanThreads / 32];
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) cross += __shfl_down_sync(0xffffffffu, cross, s);
  if ((t & 31) == 0) s_x[t >> 5] = cross;
  __syncthreads();
  if (t == 0) {
    unsigned long long x = 0;
    for (int w = 0; w < kAucScanThreads / 32; ++w) x += s_x[w];
    a.cross[blockIdx.x] = x;
    if (blockIdx.x == kAucScanBlocks - 1) out->npos = base + tot;
  }
  if (blockIdx.x == 0) {
    // (fused worker path) the evaluation partial sums of the step's blocks, in a fixed order
    __shared__ double s_o[kAucScanThreads / 32];
    __shared__ unsigned long long s_c[kAucScanThreads / 32];
    double o = 0; unsigned long long c = 0;
    for (unsigned i = t; i < nparts; i += kAucScanThreads) { o += part_objv[i]; c += part_correct[i]; }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) { o += __shfl_down_sync(0xffffffffu, o, s); c += __shfl_down_sync(0xffffffffu, c, s); }
    if ((t & 31) == 0) { s_o[t >> 5] = o; s_c[t >> 5] = c; }
    __syncthreads();
    if (t == 0) {
      double oo = 0; unsigned long long cc = 0;
      for (int w = 0; w < kAucScanThreads / 32; ++w) { oo += s_o[w]; cc += s_c[w]; }
      out->area_lo = 0; out->ties = 0; out->ngroups = 0;   // the count kernel adds into these
      if (nparts) { out->objv = oo; out->logloss = 0; out->correct = cc; }
    }
  }
}

__global__ void __launch_bounds__(256)
auc_scatter_kernel(size_t n, const float* __restrict__ label, const float* __restrict__ xw, AucScratch a,
                   unsigned short* __restrict__ pos_low) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  // the scan has read the histogram: leave it zero for the next minibatch
  for (size_t h = i; h < kAucBins; h += (size_t)gridDim.x * blockDim.x) a.hist[h] = 0ull;
  const bool pos = i < n && label[i] > 0;
  const unsigned ob = pos ? ordered_bits(xw[i]) : 0u;
  const unsigned b = ob >> 16, lo = ob & 0xffffu;
  const unsigned peers = __ballot_sync(0xffffffffu, pos);
  if (!pos) return;
  // an untrained or very sparse model gives every example the same score: one slot counter would take a ticket
  // from every positive in turn. Lanes that all sit in one bin take their tickets with one atomic.
  int same;
  __match_all_sync(peers, b, &same);
  unsigned slot;
  if (same) {
    const int leader = __ffs(peers) - 1;
    const unsigned lane = threadIdx.x & 31;
    unsigned base = 0;
    const unsigned mn = __reduce_min_sync(peers, lo), mx = __reduce_max_sync(peers, lo);
    if ((int)lane == leader) {
      base = atomicAdd(a.cursor + b, (unsigned)__popc(peers));
      atomicMin(a.lo_min + b, mn); atomicMax(a.lo_max + b, mx);
    }
    base = __shfl_sync(peers, base, leader);
    slot = base + __popc(peers & lanemask_lt());
  } else {
    slot = atomicAdd(a.cursor + b, 1u);
    atomicMin(a.lo_min + b, lo); atomicMax(a.lo_max + b, lo);
  }
  pos_low[slot] = (unsigned short)lo;
}

// Heavy bins: a block sorts the bin's positives in shared memory (bitonic, 16-bit keys), in place.
__device__ __forceinline__ bool auc_bin_sorted(unsigned n, unsigned mn, unsigned mx) {
  return n > kAucHeavy && n <= kAucSortMax && mn != mx;
}
__global__ void __launch_bounds__(1024)
auc_sort_heavy_kernel(AucScratch a, unsigned short* __restrict__ pos_low) {
  extern __shared__ unsigned short s_k[];
  const unsigned nheavy = *a.nheavy;
  for (unsigned h = blockIdx.x; h < nheavy; h += gridDim.x) {
    const unsigned b = a.heavy[h], beg = a.begin[b], n = a.cursor[b] - beg;
    if (!auc_bin_sorted(n, a.lo_min[b], a.lo_max[b])) continue;   // (uniform per block)
    unsigned N = 128;
    while (N < n) N <<= 1;
    for (unsigned i = threadIdx.x; i < N; i += blockDim.x) s_k[i] = i < n ? pos_low[beg + i] : (unsigned short)0xffffu;
    __syncthreads();
    for (unsigned k = 2; k <= N; k <<= 1)
      for (unsigned j = k >> 1; j > 0; j >>= 1) {
        for (unsigned i = threadIdx.x; i < N; i += blockDim.x) {
          const unsigned p = i ^ j;
          if (p > i) {
            const unsigned short x = s_k[i], y = s_k[p];
            if (((i & k) == 0) == (x > y)) { s_k[i] = y; s_k[p] = x; }
          }
        }
        __syncthreads();
      }
    for (unsigned i = threadIdx.x; i < n; i += blockDim.x) pos_low[beg + i] = s_k[i];
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256)
auc_count_kernel(size_t n, const float* __restrict__ label, const float* __restrict__ xw, AucScratch a,
                 const unsigned short* __restrict__ pos_low, EvalOut* out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long lt = 0, eq = 0;
  if (blockIdx.x == 0 && threadIdx.x < kAucScanBlocks) lt = a.cross[threadIdx.x];   // the pairs the bins alone settle
  if (blockIdx.x == 0 && threadIdx.x == 0) *a.nheavy = 0u;                          // (the sort kernel has finished)
  if (i < n && !(label[i] > 0)) {
    const unsigned ob = ordered_bits(xw[i]);
    const unsigned b = ob >> 16, lo = ob & 0xffffu;
    const unsigned beg = a.begin[b], end = a.cursor[b];
    if (end > beg) {
      const unsigned mn = a.lo_min[b], mx = a.lo_max[b], cnt = end - beg;
      if (mn == mx) {            // all positives of the bin share one score
        if (mn < lo) lt += cnt; else if (mn == lo) eq = cnt;
      } else if (mx < lo) {
        lt += cnt;
      } else if (mn <= lo) {
        if (auc_bin_sorted(cnt, mn, mx)) {
          unsigned l0 = beg, h0 = end;          // first position with a key >= lo
          while (l0 < h0) { const unsigned m = (l0 + h0) >> 1; if (pos_low[m] < lo) l0 = m + 1; else h0 = m; }
          unsigned l1 = l0, h1 = end;           // first position with a key > lo
          while (l1 < h1) { const unsigned m = (l1 + h1) >> 1; if (pos_low[m] <= lo) l1 = m + 1; else h1 = m; }
          lt += l0 - beg; eq = l1 - l0;
        } else {
          unsigned l = 0, e = 0;
          for (unsigned j = beg; j < end; ++j) { const unsigned p = pos_low[j]; l += p < lo; e += p == lo; }
          lt += l; eq = e;
        }
      }
    }
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) { lt += __shfl_xor_sync(0xffffffffu, lt, s); eq += __shfl_xor_sync(0xffffffffu, eq, s); }
  __shared__ unsigned long long s_l[8], s_e[8];
  if ((threadIdx.x & 31) == 0) { s_l[threadIdx.x >> 5] = lt; s_e[threadIdx.x >> 5] = eq; }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long l = 0, e = 0;
    for (int w = 0; w < 8; ++w) { l += s_l[w]; e += s_e[w]; }
    if (l) atomicAdd(&out->area_lo, l);
    if (e) atomicAdd(&out->ties, e);
  }
}

constexpr size_t kAucCountMax = 1u << 20;
int auc_scratch_get(whContext* ctx, void** base) {
  if (ctx->auc.cap < kAucScratchBytes) {
    WH_TRY(ctx->auc.reserve(kAucScratchBytes));
    WH_CUDA(cudaMemsetAsync(ctx->auc.p, 0, kAucScratchBytes, ctx->stream));
  }
  *base = ctx->auc.p;
  return WH_OK;
}
unsigned long long* auc_hist_ptr(void* scratch_base) { return static_cast<unsigned long long*>(scratch_base); }

// hist_filled: the caller's kernel has already added every example to the histogram of this context's scratch.
int auc_count_launch(whContext* ctx, size_t n, const float* label, const float* xw, EvalOut* out, bool hist_filled,
                     unsigned nparts, const double* part_objv, const unsigned* part_correct) {
  void* base;
  WH_TRY(auc_scratch_get(ctx, &base));
  AucScratch a = auc_scratch(base);
  WH_TRY(ctx->sort_tmp[0].reserve(n * 2 + 64));
  unsigned short* pos_low = ctx->sort_tmp[0].as<unsigned short>();
  const unsigned nb = (unsigned)div_up(n, 256);
  if (!hist_filled) WH_LAUNCH(auc_hist_kernel, nb, 256, 0, ctx->stream, n, label, xw, a.hist);
  WH_LAUNCH(auc_scan_kernel, kAucScanBlocks, kAucScanThreads, 0, ctx->stream, a, nparts, part_objv, part_correct, out);
  WH_LAUNCH(auc_scatter_kernel, nb, 256, 0, ctx->stream, n, label, xw, a, pos_low);
  WH_CUDA(cudaFuncSetAttribute(auc_sort_heavy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kAucSortMax * 2)));
  WH_LAUNCH(auc_sort_heavy_kernel, 16, 1024, kAucSortMax * 2, ctx->stream, a, pos_low);
  WH_LAUNCH(auc_count_kernel, nb, 256, 0, ctx->stream, n, label, xw, a, pos_low, out);
  return WH_OK;
}

// ---- AUC in ONE launch: a thread-block cluster sorts and counts -------------------------------
// 8 CTAs x 1024 threads (one cluster: hardware barrier + distributed shared memory) run the
// whole AUC: build the keys, LSD radix sort them (8-bit digits, constant digits skipped), then
// count pairs. Per pass each warp owns a contiguous run, so "rank among equal digits" is
// (earlier CTAs) + (earlier warps) + (earlier rounds) + (lower lanes); CTA totals are exchanged
// through DSMEM between two cluster barriers. The sort of 100k 33-bit keys costs ~10 cluster
// barriers instead of ~30 kernel launches.
namespace cg = cooperative_groups;
constexpr int kAucCtas = 8, kAucThreads = 1024, kAucWarps = kAucThreads / 32;
constexpr size_t kAucClusterMax = 1u << 20;

__device__ __forceinline__ uint64_t auc_key(const float* __restrict__ label, const float* __restrict__ xw, size_t i) {
  return ((uint64_t)ordered_bits(xw[i]) << 1) | (label[i] > 0 ? 1u : 0u);
}

__global__ void __cluster_dims__(kAucCtas, 1, 1) __launch_bounds__(kAucThreads)
auc_cluster_kernel(size_t n, const float* __restrict__ label, const float* __restrict__ xw,
                   uint64_t* bufA, uint64_t* bufB, uint32_t* gpos, uint32_t* gneg, EvalOut* out) {
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned cta = cluster.block_rank();
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned lt = lanemask_lt();
  __shared__ unsigned cnt[kAucWarps][256];
  __shared__ unsigned tot[256];
  __shared__ unsigned scan_tmp[8];
  __shared__ unsigned long long s_bits[2];          // OR / AND of this CTA's keys
  __shared__ unsigned w_pos[kAucWarps], w_head[kAucWarps];
  __shared__ unsigned c_pos, c_head;                // this CTA's totals
  __shared__ unsigned long long s_red[2][kAucWarps];

  // each of the 256 warps of the cluster owns a contiguous, 32-aligned run of the array
  const size_t chunk = ((n + kAucCtas * kAucWarps - 1) / (kAucCtas * kAucWarps) + 31) / 32 * 32;
  const size_t gw = (size_t)cta * kAucWarps + warp;
  const size_t lo = min(n, gw * chunk), hi = min(n, lo + chunk);

  // ---- keys + which bytes vary
  if (threadIdx.x == 0) { s_bits[0] = 0ull; s_bits[1] = ~0ull; }
  __syncthreads();
  {
    unsigned long long o = 0ull, a = ~0ull;
    for (size_t i = lo + lane; i < hi; i += 32) {
      uint64_t k = auc_key(label, xw, i);
      bufA[i] = k; o |= k; a &= k;
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) { o |= __shfl_xor_sync(0xffffffffu, o, s); a &= __shfl_xor_sync(0xffffffffu, a, s); }
    if (lane == 0) { atomicOr(&s_bits[0], o); atomicAnd(&s_bits[1], a); }
  }
  __threadfence();
  cluster.sync();
  unsigned long long diff;
  {
    unsigned long long o = 0ull, a = ~0ull;
    for (int r = 0; r < kAucCtas; ++r) {
      const unsigned long long* rb = cluster.map_shared_rank(s_bits, r);
      o |= rb[0]; a &= rb[1];
    }
    diff = o ^ a;
  }

  // ---- LSD passes over the bytes that vary (33-bit keys: bytes 0..4)
  uint64_t* src = bufA;
  uint64_t* dst = bufB;
  for (int byte = 0; byte < 5; ++byte) {
    if (((diff >> (8 * byte)) & 0xffull) == 0) continue;   // same decision in every CTA
    const int shift = 8 * byte;
    for (int w = 0; w < kAucWarps; ++w) if (threadIdx.x < 256) cnt[w][threadIdx.x] = 0;
    __syncthreads();
    for (size_t i = lo + lane; i < hi + ((32 - ((hi - lo) & 31)) & 31); i += 32) {   // full rounds
      const bool valid = i < hi;
      unsigned d = valid ? (unsigned)((src[i] >> shift) & 0xff) : 0u;
      unsigned peers = match_any_bits<8>(d, valid);
      if (valid && (__ffs(peers) - 1) == (int)lane) cnt[warp][d] += __popc(peers);
      __syncwarp();
    }
    __syncthreads();
    if (threadIdx.x < 256) {
      unsigned run = 0;
      for (int w = 0; w < kAucWarps; ++w) { unsigned c = cnt[w][threadIdx.x]; cnt[w][threadIdx.x] = run; run += c; }
      tot[threadIdx.x] = run;
    }
    cluster.sync();
    {
      unsigned before = 0, total = 0;
      if (threadIdx.x < 256) {
        for (int r = 0; r < kAucCtas; ++r) {
          unsigned v = cluster.map_shared_rank(tot, r)[threadIdx.x];
          if (r < (int)cta) before += v;
          total += v;
        }
      }
      // exclusive scan of `total` over the 256 digits (warps 0..7)
      unsigned incl = total;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { unsigned t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
      if (warp < 8 && lane == 31) scan_tmp[warp] = incl;
      __syncthreads();
      if (threadIdx.x < 256) {
        unsigned wbase = 0;
        for (int w = 0; w < (int)warp; ++w) wbase += scan_tmp[w];
        unsigned base = wbase + incl - total + before;
        for (int w = 0; w < kAucWarps; ++w) cnt[w][threadIdx.x] += base;
      }
    }
    __syncthreads();
    for (size_t i = lo + lane; i < hi + ((32 - ((hi - lo) & 31)) & 31); i += 32) {
      const bool valid = i < hi;
      uint64_t k = valid ? src[i] : 0;
      unsigned d = valid ? (unsigned)((k >> shift) & 0xff) : 0u;
      unsigned peers = match_any_bits<8>(d, valid);
      unsigned old = valid ? cnt[warp][d] : 0;
      __syncwarp();
      if (valid && (__ffs(peers) - 1) == (int)lane) cnt[warp][d] = old + __popc(peers);
      __syncwarp();
      if (valid) dst[old + __popc(peers & lt)] = k;
    }
    __threadfence();
    cluster.sync();   // dst complete and visible; `tot` free for reuse
    uint64_t* t = src; src = dst; dst = t;
  }
  const uint64_t* sorted = src;

  // ---- pair counting. Sorted ascending by score, negatives first among equal scores.
  // sweep 1: per-warp number of positives and of score-group heads
  unsigned npos = 0, nhead = 0;
  for (size_t i = lo + lane; i < hi + ((32 - ((hi - lo) & 31)) & 31); i += 32) {
    const bool valid = i < hi;
    uint64_t k = valid ? sorted[i] : 0;
    uint64_t prev = (valid && i > 0) ? sorted[i - 1] : ~0ull;
    bool head = valid && (i == 0 || (k >> 1) != (prev >> 1));
    npos += __popc(__ballot_sync(0xffffffffu, valid && (k & 1)));
    nhead += __popc(__ballot_sync(0xffffffffu, head));
  }
  if (lane == 0) { w_pos[warp] = npos; w_head[warp] = nhead; }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned p = 0, h = 0;
    for (int w = 0; w < kAucWarps; ++w) { unsigned a = w_pos[w], b = w_head[w]; w_pos[w] = p; w_head[w] = h; p += a; h += b; }
    c_pos = p; c_head = h;
  }
  cluster.sync();
  unsigned pos_before = w_pos[warp], head_before = w_head[warp], P = 0, G = 0;
  for (int r = 0; r < kAucCtas; ++r) {
    unsigned p = *cluster.map_shared_rank(&c_pos, r), h = *cluster.map_shared_rank(&c_head, r);
    if (r < (int)cta) { pos_before += p; head_before += h; }
    P += p; G += h;
  }
  // sweep 2: at every group head remember how many positives / negatives precede it
  for (size_t i = lo + lane; i < hi + ((32 - ((hi - lo) & 31)) & 31); i += 32) {
    const bool valid = i < hi;
    uint64_t k = valid ? sorted[i] : 0;
    uint64_t prev = (valid && i > 0) ? sorted[i - 1] : ~0ull;
    bool head = valid && (i == 0 || (k >> 1) != (prev >> 1));
    unsigned pb = __ballot_sync(0xffffffffu, valid && (k & 1)), hb = __ballot_sync(0xffffffffu, head);
    if (head) {
      unsigned g = head_before + __popc(hb & lt), pbefore = pos_before + __popc(pb & lt);
      gpos[g] = pbefore;
      gneg[g] = (unsigned)i - pbefore;
    }
    pos_before += __popc(pb); head_before += __popc(hb);
  }
  __threadfence();
  cluster.sync();
  // per group g: its negatives each see gpos[g] positives strictly below; ties = P_g * N_g
  const unsigned N = (unsigned)n - P;
  unsigned long long area = 0, ties = 0;
  for (size_t g = (size_t)cta * kAucThreads + threadIdx.x; g < G; g += (size_t)kAucCtas * kAucThreads) {
    unsigned p0 = gpos[g], n0 = gneg[g];
    unsigned p1 = (g + 1 < G) ? gpos[g + 1] : P, n1 = (g + 1 < G) ? gneg[g + 1] : N;
    area += (unsigned long long)(n1 - n0) * p0;
    ties += (unsigned long long)(n1 - n0) * (p1 - p0);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { area += __shfl_xor_sync(0xffffffffu, area, o); ties += __shfl_xor_sync(0xffffffffu, ties, o); }
  if (lane == 0) { s_red[0][warp] = area; s_red[1][warp] = ties; }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long a = 0, t = 0;
    for (int w = 0; w < kAucWarps; ++w) { a += s_red[0][w]; t += s_red[1][w]; }
    if (a) atomicAdd(&out->area_lo, a);
    if (t) atomicAdd(&out->ties, t);
    if (cta == 0) { out->npos = P; out->ngroups = G; }
  }
  cluster.sync();   // no CTA may exit while a peer can still read its shared memory
}


// Register-resident variant for the usual minibatch sizes (n <= 128 warps x 32 x R): every warp
// keeps its run of keys in registers, so a pass costs one L2 round trip for the loads, one for
// the scatter, and three cluster barriers.
constexpr int kAucRegThreads = 512, kAucRegWarps = kAucRegThreads / 32;

template <int R>
__global__ void __cluster_dims__(kAucCtas, 1, 1) __launch_bounds__(kAucRegThreads)
auc_cluster_reg_kernel(size_t n, const float* __restrict__ label, const float* __restrict__ xw,
                       uint64_t* bufA, uint64_t* bufB, uint32_t* gpos, uint32_t* gneg, EvalOut* out) {
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned cta = cluster.block_rank();
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned lt = lanemask_lt();
  __shared__ unsigned cnt[kAucRegWarps][256];
  __shared__ unsigned tot[256];
  __shared__ unsigned scan_tmp[8];
  __shared__ unsigned long long s_bits[2];
  __shared__ unsigned w_pos[kAucRegWarps], w_head[kAucRegWarps];
  __shared__ unsigned c_pos, c_head;
  __shared__ unsigned long long s_red[2][kAucRegWarps];

  const size_t chunk = 32 * R;
  const size_t gw = (size_t)cta * kAucRegWarps + warp;
  const size_t lo = min(n, gw * chunk), hi = min(n, lo + chunk);

  uint64_t k[R];
  if (threadIdx.x == 0) { s_bits[0] = 0ull; s_bits[1] = ~0ull; }
  __syncthreads();
  {
    unsigned long long o = 0ull, a = ~0ull;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const size_t i = lo + r * 32 + lane;
      k[r] = 0;
      if (i < hi) { k[r] = auc_key(label, xw, i); o |= k[r]; a &= k[r]; }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) { o |= __shfl_xor_sync(0xffffffffu, o, s); a &= __shfl_xor_sync(0xffffffffu, a, s); }
    if (lane == 0) { atomicOr(&s_bits[0], o); atomicAnd(&s_bits[1], a); }
  }
  cluster.sync();
  unsigned long long diff;
  {
    unsigned long long o = 0ull, a = ~0ull;
    for (int r = 0; r < kAucCtas; ++r) {
      const unsigned long long* rb = cluster.map_shared_rank(s_bits, r);
      o |= rb[0]; a &= rb[1];
    }
    diff = o ^ a;
  }

  uint64_t* dst = bufA;
  uint64_t* other = bufB;
  bool in_regs = true;   // before the first active pass the keys exist only in registers
  for (int byte = 0; byte < 5; ++byte) {
    if (((diff >> (8 * byte)) & 0xffull) == 0) continue;
    const int shift = 8 * byte;
    if (!in_regs) {
#pragma unroll
      for (int r = 0; r < R; ++r) { const size_t i = lo + r * 32 + lane; k[r] = (i < hi) ? other[i] : 0; }
    }
    for (int w = 0; w < kAucRegWarps; ++w) if (threadIdx.x < 256) cnt[w][threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const bool valid = lo + r * 32 + lane < hi;
      unsigned d = valid ? (unsigned)((k[r] >> shift) & 0xff) : 0u;
      unsigned peers = match_any_bits<8>(d, valid);
      if (valid && (__ffs(peers) - 1) == (int)lane) cnt[warp][d] += __popc(peers);
      __syncwarp();
    }
    __syncthreads();
    if (threadIdx.x < 256) {
      unsigned run = 0;
#pragma unroll
      for (int w = 0; w < kAucRegWarps; ++w) { unsigned c = cnt[w][threadIdx.x]; cnt[w][threadIdx.x] = run; run += c; }
      tot[threadIdx.x] = run;
    }
    cluster.sync();
    {
      unsigned before = 0, total = 0;
      if (threadIdx.x < 256) {
#pragma unroll
        for (int r = 0; r < kAucCtas; ++r) {
          unsigned v = cluster.map_shared_rank(tot, r)[threadIdx.x];
          if (r < (int)cta) before += v;
          total += v;
        }
      }
      unsigned incl = total;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { unsigned t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
      if (warp < 8 && lane == 31) scan_tmp[warp] = incl;
      __syncthreads();
      if (threadIdx.x < 256) {
        unsigned wbase = 0;
        for (int w = 0; w < (int)warp; ++w) wbase += scan_tmp[w];
        unsigned base = wbase + incl - total + before;
#pragma unroll
        for (int w = 0; w < kAucRegWarps; ++w) cnt[w][threadIdx.x] += base;
      }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const bool valid = lo + r * 32 + lane < hi;
      unsigned d = valid ? (unsigned)((k[r] >> shift) & 0xff) : 0u;
      unsigned peers = match_any_bits<8>(d, valid);
      unsigned old = valid ? cnt[warp][d] : 0;
      __syncwarp();
      if (valid && (__ffs(peers) - 1) == (int)lane) cnt[warp][d] = old + __popc(peers);
      __syncwarp();
      if (valid) dst[old + __popc(peers & lt)] = k[r];
    }
    __threadfence();
    cluster.sync();
    uint64_t* t = dst; dst = other; other = t;   // `other` now holds the pass's output
    in_regs = false;
  }
  if (!in_regs) {
#pragma unroll
    for (int r = 0; r < R; ++r) { const size_t i = lo + r * 32 + lane; k[r] = (i < hi) ? other[i] : 0; }
  }
  // the key just before this warp's run (all keys equal if no pass ran: then nothing is a head but i == 0)
  uint64_t before_run = (lo > 0 && lo < n) ? (in_regs ? k[0] : other[lo - 1]) : ~0ull;
  if (in_regs && lo > 0) before_run = __shfl_sync(0xffffffffu, k[0], 0);

  unsigned npos = 0, nhead = 0;
  unsigned posb[R], headb[R];
  uint64_t carry = before_run;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const size_t i = lo + r * 32 + lane;
    const bool valid = i < hi;
    uint64_t prev = __shfl_up_sync(0xffffffffu, k[r], 1);
    if (lane == 0) prev = carry;
    carry = __shfl_sync(0xffffffffu, k[r], 31);
    bool head = valid && (i == 0 || (k[r] >> 1) != (prev >> 1));
    posb[r] = __ballot_sync(0xffffffffu, valid && (k[r] & 1));
    headb[r] = __ballot_sync(0xffffffffu, head);
    npos += __popc(posb[r]); nhead += __popc(headb[r]);
  }
  if (lane == 0) { w_pos[warp] = npos; w_head[warp] = nhead; }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned p = 0, h = 0;
    for (int w = 0; w < kAucRegWarps; ++w) { unsigned a = w_pos[w], b = w_head[w]; w_pos[w] = p; w_head[w] = h; p += a; h += b; }
    c_pos = p; c_head = h;
  }
  cluster.sync();
  unsigned pos_before = w_pos[warp], head_before = w_head[warp], P = 0, G = 0;
  for (int r = 0; r < kAucCtas; ++r) {
    unsigned p = *cluster.map_shared_rank(&c_pos, r), h = *cluster.map_shared_rank(&c_head, r);
    if (r < (int)cta) { pos_before += p; head_before += h; }
    P += p; G += h;
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const size_t i = lo + r * 32 + lane;
    if ((headb[r] >> lane) & 1u) {
      unsigned g = head_before + __popc(headb[r] & lt), pbefore = pos_before + __popc(posb[r] & lt);
      gpos[g] = pbefore;
      gneg[g] = (unsigned)i - pbefore;
    }
    pos_before += __popc(posb[r]); head_before += __popc(headb[r]);
  }
  __threadfence();
  cluster.sync();
  const unsigned N = (unsigned)n - P;
  unsigned long long area = 0, ties = 0;
  for (size_t g = (size_t)cta * kAucRegThreads + threadIdx.x; g < G; g += (size_t)kAucCtas * kAucRegThreads) {
    unsigned p0 = gpos[g], n0 = gneg[g];
    unsigned p1 = (g + 1 < G) ? gpos[g + 1] : P, n1 = (g + 1 < G) ? gneg[g + 1] : N;
    area += (unsigned long long)(n1 - n0) * p0;
    ties += (unsigned long long)(n1 - n0) * (p1 - p0);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { area += __shfl_xor_sync(0xffffffffu, area, o); ties += __shfl_xor_sync(0xffffffffu, ties, o); }
  if (lane == 0) { s_red[0][warp] = area; s_red[1][warp] = ties; }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long a = 0, t = 0;
    for (int w = 0; w < kAucRegWarps; ++w) { a += s_red[0][w]; t += s_red[1][w]; }
    if (a) atomicAdd(&out->area_lo, a);
    if (t) atomicAdd(&out->ties, t);
    if (cta == 0) { out->npos = P; out->ngroups = G; }
  }
  cluster.sync();
}

int auc_launch(whContext* ctx, size_t n, const float* label, const float* xw, EvalOut* out, bool few_sms) {
  if (n == 0) return WH_OK;
  WH_REQUIRE(n < (size_t)INT32_MAX, "AUC: %zu examples exceed 32-bit counters", n);
  // Two ways to the same integers. Counting: five short launches over the whole GPU (47 us in all). One-cluster sort:
  // one launch of 165 us on 8 SMs. The second is the cheaper one in GPU time, and where the step's own kernels fill the
  // GPU back to back (sharded jobs with several minibatches in flight: `few_sms`) it is the faster one for the job --
  // measured on 2 GPUs: 0.161 against 0.172 ms per step. WH_AUC=cluster / WH_AUC=count force one.
  static int forced = -1;
  if (forced < 0) { const char* e = getenv("WH_AUC"); forced = !e ? 0 : (e[0] == 'c' && e[1] == 'l') ? 1 : 2; }
  const bool cluster = forced == 1 || (forced == 0 && few_sms);
  if (!cluster && n <= kAucCountMax) return auc_count_launch(ctx, n, label, xw, out, false, 0, nullptr, nullptr);
  if (n <= kAucClusterMax) {
    WH_TRY(ctx->sort_tmp[0].reserve(n * 8)); WH_TRY(ctx->sort_tmp[1].reserve(n * 8));
    WH_TRY(ctx->sort_tmp[5].reserve((n + 1) * 8));
    uint32_t* gpos = ctx->sort_tmp[5].as<uint32_t>();
    const size_t warps = (size_t)kAucCtas * kAucRegWarps;
#define WH_AUC_REG(R)                                                                                     \
    if (n <= warps * 32 * (R)) {                                                                           \
      WH_LAUNCH(auc_cluster_reg_kernel<R>, kAucCtas, kAucRegThreads, 0, ctx->stream, n, label, xw,         \
                ctx->sort_tmp[0].as<uint64_t>(), ctx->sort_tmp[1].as<uint64_t>(), gpos, gpos + n + 1, out); \
      return WH_OK;                                                                                        \
    }
    WH_AUC_REG(4) WH_AUC_REG(8) WH_AUC_REG(16) WH_AUC_REG(32)
#undef WH_AUC_REG
    WH_LAUNCH(auc_cluster_kernel, kAucCtas, kAucThreads, 0, ctx->stream, n, label, xw, ctx->sort_tmp[0].as<uint64_t>(),
              ctx->sort_tmp[1].as<uint64_t>(), gpos, gpos + n + 1, out);
    return WH_OK;
  }
  return auc_launch_multiblock(ctx, n, label, xw, out);
}

EvalHost eval_finish(const EvalOut& e, size_t rows) {
  EvalHost h;
  h.objv = e.objv;
  h.logloss = -e.logloss;
  // binary_class_evaluation.h:49-50
  float acc = rows ? (float)e.correct / (float)rows : 0.f;
  h.acc = acc > 0.5f ? acc : 1.f - acc;
  // binary_class_evaluation.h:36-38
  const double P = (double)e.npos, N = (double)rows - P;
  if (e.npos == 0 || e.npos == rows) { h.auc = h.auc_lo = h.auc_hi = 1.0; return h; }
  const double lo = (double)e.area_lo / (P * N), hi = (double)(e.area_lo + e.ties) / (P * N);
  auto fold = [](double a) { return a < 0.5 ? 1.0 - a : a; };
  h.auc = fold(0.5 * (lo + hi));
  const double flo = fold(lo), fhi = fold(hi);
  h.auc_hi = flo > fhi ? flo : fhi;
  h.auc_lo = (lo <= 0.5 && hi >= 0.5) ? 0.5 : (flo < fhi ? flo : fhi);
  return h;
}

}  // namespace wh

using namespace wh;

extern "C" {

int wh_bin_eval(whContext* ctx, const float* label, const float* predict, size_t n, int mem, float* out6) {
  WH_REQUIRE(ctx && out6 && (n == 0 || (label && predict)), "wh_bin_eval: NULL argument");
  DeviceGuard g(ctx->device);
  const float *dl, *dp;
  WH_TRY(to_device(ctx, ctx->stage[0], label, n, mem, &dl));
  WH_TRY(to_device(ctx, ctx->stage[1], predict, n, mem, &dp));
  EvalOut* dev = reinterpret_cast<EvalOut*>(ctx->small.as<char>() + 4096);
  WH_TRY(loss_forward_launch(ctx, WH_LOSS_LOGIT, n, dl, dp, nullptr, dev, 1));
  WH_TRY(auc_launch(ctx, n, dl, dp, dev));
  WH_CUDA(cudaMemcpyAsync(ctx->pinned.p, dev, sizeof(EvalOut), cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  EvalHost h = eval_finish(*ctx->pinned.as<EvalOut>(), n);
  out6[0] = (float)h.auc; out6[1] = (float)h.acc; out6[2] = (float)h.logloss; out6[3] = (float)h.objv;
  out6[4] = (float)h.auc_lo; out6[5] = (float)h.auc_hi;
  return WH_OK;
}

int wh_linear_loss(whContext* ctx, int loss, const whLocalBlock* D, int mem, const float* w, size_t U,
                   double* prog, float* grad, float* xw, float* dual) {
  WH_REQUIRE(ctx && D && prog, "wh_linear_loss: NULL argument");
  WH_REQUIRE(loss == WH_LOSS_LOGIT || loss == WH_LOSS_SQUARE_HINGE || loss == WH_LOSS_SQUARE, "wh_linear_loss: unknown loss type %d", loss);
  DeviceGuard g(ctx->device);
  DevLocal d;
  WH_TRY(stage_local(ctx, D, mem, &d));
  WH_REQUIRE(d.rows == 0 || d.label != nullptr, "wh_linear_loss: label is NULL");
  const float* dw; float *dxw, *ddual, *dgrad = nullptr;
  WH_TRY(to_device(ctx, ctx->stage[4], w, U, mem, &dw));
  // Init: Xw = X w (loss.h:32-39); Xw_ starts zeroed (vector::resize), rows without entries stay 0
  WH_TRY(out_buffer(ctx, ctx->stage[5], xw, d.rows, xw ? mem : WH_HOST, &dxw));
  WH_TRY(out_buffer(ctx, ctx->stage[6], dual, d.rows, dual ? mem : WH_HOST, &ddual));
  if (d.rows) WH_CUDA(cudaMemsetAsync(dxw, 0, d.rows * 4, ctx->stream));
  WH_TRY(spmv_times_launch(ctx, d.rows, d.offset, d.index, d.value, dw, dxw));
  EvalOut* dev = reinterpret_cast<EvalOut*>(ctx->small.as<char>() + 4096);
  WH_TRY(loss_forward_launch(ctx, loss, d.rows, d.label, dxw, ddual, dev, 0));
  WH_TRY(auc_launch(ctx, d.rows, d.label, dxw, dev));
  if (grad) {  // CalcGrad: grad = X^T dual (loss.h:114), squared hinge then scales by -2 (:152-155)
    long long* acc;
    WH_TRY(ensure_acc(ctx, U, &acc));
    FixedScale* fs = reinterpret_cast<FixedScale*>(ctx->small.as<char>() + 1024);
    // the staged w is dead after Init: reuse its slot for the gradient of host callers
    WH_TRY(out_buffer(ctx, ctx->stage[4], grad, U, mem, &dgrad));
    if (d.rows == 0) { if (U) WH_CUDA(cudaMemsetAsync(dgrad, 0, U * 4, ctx->stream)); }
    else WH_TRY(spmv_trans_times_launch(ctx, d.rows, d.nnz, d.offset, d.index, d.value, ddual, dgrad, U, acc, fs,
                                        loss == WH_LOSS_SQUARE_HINGE ? -2.f : 1.f, acc_copies()));
    WH_TRY(out_commit(ctx, grad, dgrad, U, mem));
  }
  if (xw) WH_TRY(out_commit(ctx, xw, dxw, d.rows, mem));
  if (dual) WH_TRY(out_commit(ctx, dual, ddual, d.rows, mem));
  WH_CUDA(cudaMemcpyAsync(ctx->pinned.p, dev, sizeof(EvalOut), cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  EvalHost h = eval_finish(*ctx->pinned.as<EvalOut>(), d.rows);
  // learn/linear/progress.h:28-34
  prog[0] = h.objv; prog[1] = h.acc; prog[2] = h.auc; prog[3] = 1; prog[4] = (double)d.rows; prog[5] = 0;
  return WH_OK;
}

}  // extern "C"
