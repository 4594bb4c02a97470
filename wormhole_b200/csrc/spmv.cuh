// Pieces (2) and (5): CSR products on a localized block. Internal launch API (device pointers).
//
// Forward products (Times) walk each row sequentially in CSR order with one thread per row (or
// one thread per row and column for SpMM), which reproduces the reference's fp32 summation
// order exactly (learn/base/spmv.h:72-91, learn/base/spmm.h:56-81). The block first stages its
// rows' index/value run in shared memory with coalesced loads, so global traffic is one
// streaming pass over the CSR plus the gathers.
//
// Transposed products (TransTimes) cannot keep the reference's order (ascending rows per output,
// learn/base/spmv.h:103-117) without serialising hot features, so they accumulate in 64-bit
// FIXED POINT: every term x_i*val_ij is computed in fp32 exactly as the reference does, scaled
// by a power of two chosen from max|x|, max|val| and the entry count, and added with integer
// atomics. Integer addition is associative, so the result is independent of thread order
// (bit-reproducible run to run) and differs from the reference only by the reference's own
// fp32 accumulation rounding (tests bound it by 1e-5 of the sum of |terms|).
#pragma once
#include <math.h>

#include "common.cuh"

namespace wh {

// replicas of the accumulator array that blocks spread their atomics over (power of two)
constexpr unsigned kAccCopies = 8;   // upper bound; acc_copies() picks the count in use
unsigned acc_copies();
#ifdef __CUDACC__
// Sum of one output's accumulator replicas, clearing them for the next minibatch. All replicas are loaded
// first, then cleared: interleaved, every load waits for the one before it (the stores may alias) and a
// thread pays `copies` memory latencies in a row.
__device__ __forceinline__ long long sum_and_clear_replicas(long long* acc, unsigned copies, size_t stride, size_t i) {
  long long v[kAccCopies];
#pragma unroll
  for (unsigned c = 0; c < kAccCopies; ++c) v[c] = c < copies ? __ldcg(acc + c * stride + i) : 0;
  long long a = 0;
#pragma unroll
  for (unsigned c = 0; c < kAccCopies; ++c) a += v[c];
#pragma unroll
  for (unsigned c = 0; c < kAccCopies; ++c) if (c < copies) acc[c * stride + i] = 0;
  return a;
}
#endif

struct FixedScale {
  float scale;     // 2^e
  double inv;      // 2^-e
  unsigned maxx_bits, maxv_bits;  // float bit patterns of max|x|, max|val| (atomicMax targets)
};

// Fixed-point scale for terms bounded by 1 in magnitude: nnz * 1 * 2^e < 2^62 (the rule of fixed_scale_kernel,
// evaluated on the host when the bound is known up front: binary features, |dual| <= 1).
static inline void unit_terms_scale(size_t nnz, float* scale, double* inv) {
  int en = 0;
  while (en < 64 && ((unsigned long long)(nnz | 1) >> en)) ++en;
  const int e = 61 - 1 - en;
  *scale = ldexpf(1.f, e);
  *inv = ldexp(1.0, -e);
}
struct HotView;
struct EvalOut;
int hot_w_launch(whContext* ctx, const HotView& hot, const float* w);
int spmv_fused_hot_launch(whContext* ctx, int loss, size_t rows, size_t nnz, const size_t* offset, const uint32_t* codes,
                          const float* label, const float* w, const HotView& hot, float* xw, float* dual, EvalOut* out,
                          FixedScale* fs, long long* acc, unsigned copies, size_t stride,
                          const void* directory,   // the localizer's directory of this minibatch (uint2 words), left untouched
                          whContext* reduce_ctx = nullptr);                  // where the evaluation partials are folded (default: ctx)

int spmv_times_launch(whContext* ctx, size_t rows, const size_t* offset, const uint32_t* index,
                      const float* value, const float* x, float* y);
// acc: int64[copies * ysize], all zero on entry and left all zero on exit. fs: device scratch.
// mul: exact post-scale of the result (the -2 of the squared hinge, learn/linear/loss.h:152-155).
int spmv_trans_times_launch(whContext* ctx, size_t rows, size_t nnz, const size_t* offset,
                            const uint32_t* index, const float* value, const float* x, float* y,
                            size_t ysize, long long* acc, FixedScale* fs, float mul, unsigned copies);
int spmm_times_launch(whContext* ctx, size_t rows, const size_t* offset, const uint32_t* index,
                      const float* value, const float* x, float* y, int dim);
int spmm_trans_times_launch(whContext* ctx, size_t rows, size_t nnz, const size_t* offset,
                            const uint32_t* index, const float* value, const float* x, float p,
                            const float* z, float* y, size_t ysize, int dim, long long* acc,
                            FixedScale* fs);
// Chooses the fixed-point scale for terms x * value^vpow (max|x|, max|value| reductions + one tiny kernel).
int fixed_prepare(whContext* ctx, FixedScale* fs, const float* x, size_t nx, const float* value, size_t nnz, int vpow);
int fixed_prepare_unit_x(whContext* ctx, FixedScale* fs, const float* value, size_t nnz);   // ... with |x| <= 1 known
// A whLocalBlock resolved to device pointers.
struct DevLocal {
  size_t rows = 0, nnz = 0;
  const size_t* offset = nullptr; const uint32_t* index = nullptr; const float* value = nullptr;
  const float* label = nullptr;
};
int stage_local(whContext* ctx, const whLocalBlock* D, int mem, DevLocal* out);
// zero-on-growth accumulator of the context: kAccCopies * n entries
int ensure_acc(whContext* ctx, size_t n, long long** out);

#ifdef __CUDACC__
__device__ __forceinline__ void fixed_add(long long* acc, float term, float scale) {
  long long q = __float2ll_rn(term * scale);
  atomicAdd(reinterpret_cast<unsigned long long*>(acc), static_cast<unsigned long long>(q));
}
// max|v| over an array as an atomicMax on the bit pattern (non-negative floats order like uints)
__device__ __forceinline__ void absmax_commit(unsigned* target, float m) {
  m = fabsf(m);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(target, __float_as_uint(m));
}
#endif

}  // namespace wh
