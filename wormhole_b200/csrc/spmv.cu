// Pieces (2) and (5): SpMV::Times / TransTimes and SpMM::Times / TransTimes. See spmv.cuh.
#include <stdlib.h>

#include <utility>
#include <vector>

#include "hot.cuh"
#include "loss.cuh"
#include "spmv.cuh"

namespace wh {

constexpr int kRowsPerBlock = 128;   // one thread per row
constexpr int kTileCap = 6144;       // staged entries per block (Criteo shape: 128 x 39 = 4992)

// Coalesced copy of a run of 32-bit words into shared memory with kStageIlp loads in flight per thread. (A plain
// "s[i] = g[i]" loop keeps ONE: every store waits for its load, and a thread that stages 39 words pays 39 memory
// latencies in a row -- measured as 31 % of all stall samples of the fused kernel, on this one store.)
constexpr int kStageIlp = 8;
template <typename T>
__device__ __forceinline__ void stage_run(T* s, const T* __restrict__ g, unsigned cnt) {
  for (unsigned i0 = threadIdx.x; i0 < cnt; i0 += blockDim.x * kStageIlp) {
    T v[kStageIlp];
#pragma unroll
    for (int u = 0; u < kStageIlp; ++u) { const unsigned i = i0 + u * blockDim.x; if (i < cnt) v[u] = __ldcs(g + i); }
#pragma unroll
    for (int u = 0; u < kStageIlp; ++u) { const unsigned i = i0 + u * blockDim.x; if (i < cnt) s[i] = v[u]; }
  }
}

// Loads the offsets of this block's rows and, if the rows' entries fit, their index (and value)
// run into shared memory with coalesced loads. Returns whether the run was staged.
template <bool HAS_VAL>
__device__ __forceinline__ bool stage_rows(size_t rows, const size_t* __restrict__ offset,
                                           const uint32_t* __restrict__ index,
                                           const float* __restrict__ value, size_t* s_off,
                                           uint32_t* s_idx, float* s_val) {
  const size_t r0 = (size_t)blockIdx.x * kRowsPerBlock;
  for (int t = threadIdx.x; t <= kRowsPerBlock; t += blockDim.x) s_off[t] = offset[min(r0 + t, rows)];
  __syncthreads();
  const size_t begin = s_off[0];
  const size_t cnt = s_off[kRowsPerBlock] - begin;
  const bool staged = cnt <= (size_t)kTileCap;
  if (staged) {
    stage_run(s_idx, index + begin, (unsigned)cnt);
    if (HAS_VAL) stage_run(s_val, value + begin, (unsigned)cnt);
  }
  __syncthreads();
  return staged;
}

// learn/base/spmv.h:72-91. y[i] = sum over the row, sequential fp32 in CSR order.
template <bool HAS_VAL>
__global__ void __launch_bounds__(kRowsPerBlock)
spmv_times_kernel(size_t rows, const size_t* __restrict__ offset, const uint32_t* __restrict__ index,
                  const float* __restrict__ value, const float* __restrict__ x, float* __restrict__ y) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ size_t s_off[kRowsPerBlock + 1];
  uint32_t* s_idx = reinterpret_cast<uint32_t*>(smem);
  float* s_val = reinterpret_cast<float*>(smem + sizeof(uint32_t) * kTileCap);
  const bool staged = stage_rows<HAS_VAL>(rows, offset, index, value, s_off, s_idx, s_val);
  const size_t r = (size_t)blockIdx.x * kRowsPerBlock + threadIdx.x;
  if (r >= rows) return;
  const size_t b = s_off[threadIdx.x], e = s_off[threadIdx.x + 1];
  if (b == e) return;  // spmv.h:79: rows without entries leave y untouched
  float acc = 0.f;
  if (staged) {
    const unsigned lb = (unsigned)(b - s_off[0]), le = (unsigned)(e - s_off[0]);
    unsigned j = lb;
    for (; j + 4 <= le; j += 4) {  // 4 gathers in flight, adds stay in order
      float g0 = __ldg(x + s_idx[j]), g1 = __ldg(x + s_idx[j + 1]);
      float g2 = __ldg(x + s_idx[j + 2]), g3 = __ldg(x + s_idx[j + 3]);
      if (HAS_VAL) { g0 *= s_val[j]; g1 *= s_val[j + 1]; g2 *= s_val[j + 2]; g3 *= s_val[j + 3]; }
      acc += g0; acc += g1; acc += g2; acc += g3;
    }
    for (; j < le; ++j) { float g = __ldg(x + s_idx[j]); if (HAS_VAL) g *= s_val[j]; acc += g; }
  } else {
    for (size_t j = b; j < e; ++j) { float g = __ldg(x + index[j]); if (HAS_VAL) g *= value[j]; acc += g; }
  }
  y[r] = acc;
}


// SpMV::Times with the loss as its epilogue (learn/linear/loss.h:32-39 Init, then the per-example part of
// Evaluate and CalcGrad): the thread that has just summed row i turns Xw[i] into dual[i] and adds the row's
// objective term and accuracy flag to the block's partial sums. Same summation order as spmv_times_kernel.
// HOT: index holds codes (hot.cuh): kHotFlag | slot reads the slot-indexed weight array (32 KB, L1 resident).
template <bool HOT>
__device__ __forceinline__ float gather_w(const float* __restrict__ x, const float* __restrict__ hot_w, uint32_t code) {
  // one load from a selected address: no divergent branch around two loads
  const float* p = (HOT && (code & kHotFlag)) ? hot_w + (code & (kHotSlots - 1)) : x + code;
  return __ldg(p);
}

template <bool HAS_VAL, bool HOT>
__global__ void __launch_bounds__(kRowsPerBlock)
spmv_loss_kernel(int loss, size_t rows, const size_t* __restrict__ offset, const uint32_t* __restrict__ index,
                 const float* __restrict__ value, const float* __restrict__ label, const float* __restrict__ x,
                 float* __restrict__ y, float* __restrict__ dual, double* part_objv, unsigned* part_correct,
                 const float* __restrict__ hot_w) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ size_t s_off[kRowsPerBlock + 1];
  __shared__ double s_o[kRowsPerBlock / 32];
  __shared__ unsigned s_c[kRowsPerBlock / 32];
  uint32_t* s_idx = reinterpret_cast<uint32_t*>(smem);
  float* s_val = reinterpret_cast<float*>(smem + sizeof(uint32_t) * kTileCap);
  const bool staged = stage_rows<HAS_VAL>(rows, offset, index, value, s_off, s_idx, s_val);
  const size_t r = (size_t)blockIdx.x * kRowsPerBlock + threadIdx.x;
  double objv = 0;
  unsigned correct = 0;
  if (r < rows) {
    const size_t b = s_off[threadIdx.x], e = s_off[threadIdx.x + 1];
    float acc = 0.f;
    if (staged) {
      const unsigned lb = (unsigned)(b - s_off[0]), le = (unsigned)(e - s_off[0]);
      unsigned j = lb;
      for (; j + 4 <= le; j += 4) {
        float g0 = gather_w<HOT>(x, hot_w, s_idx[j]), g1 = gather_w<HOT>(x, hot_w, s_idx[j + 1]);
        float g2 = gather_w<HOT>(x, hot_w, s_idx[j + 2]), g3 = gather_w<HOT>(x, hot_w, s_idx[j + 3]);
        if (HAS_VAL) { g0 *= s_val[j]; g1 *= s_val[j + 1]; g2 *= s_val[j + 2]; g3 *= s_val[j + 3]; }
        acc += g0; acc += g1; acc += g2; acc += g3;
      }
      for (; j < le; ++j) { float g = gather_w<HOT>(x, hot_w, s_idx[j]); if (HAS_VAL) g *= s_val[j]; acc += g; }
    } else {
      for (size_t j = b; j < e; ++j) { float g = gather_w<HOT>(x, hot_w, index[j]); if (HAS_VAL) g *= value[j]; acc += g; }
    }
    y[r] = acc;   // rows without entries: Xw_ starts zeroed in ScalarLoss::Init (vector::resize)
    float dl;
    loss_row(loss, label[r], acc, &dl, &objv, &correct);
    if (dual) dual[r] = dl;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    objv += __shfl_down_sync(0xffffffffu, objv, o);
    correct += __shfl_down_sync(0xffffffffu, correct, o);
  }
  if ((threadIdx.x & 31) == 0) { s_o[threadIdx.x >> 5] = objv; s_c[threadIdx.x >> 5] = correct; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double o = 0; unsigned c = 0;
    for (int w = 0; w < kRowsPerBlock / 32; ++w) { o += s_o[w]; c += s_c[w]; }
    part_objv[blockIdx.x] = o; part_correct[blockIdx.x] = c;
  }
}

int spmv_loss_launch(whContext* ctx, int loss, size_t rows, size_t nnz, const size_t* offset, const uint32_t* index,
                     const float* value, const float* label, const float* w, float* xw, float* dual, EvalOut* out,
                     FixedScale* fs, const float* hot_w) {
  if (rows == 0) return loss_reduce_launch(ctx, 0, nullptr, nullptr, nullptr, out, fs, nnz);
  const unsigned grid = (unsigned)div_up(rows, kRowsPerBlock);
  WH_TRY(ctx->stage[7].reserve((size_t)grid * 16 + 64));
  double* po = ctx->stage[7].as<double>();
  unsigned* pc = reinterpret_cast<unsigned*>(po + grid);
  if (value) {
    const size_t smem = kTileCap * 8;
    WH_CUDA(cudaFuncSetAttribute(spmv_loss_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    WH_CUDA(cudaFuncSetAttribute(spmv_loss_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (hot_w) WH_LAUNCH((spmv_loss_kernel<true, true>), grid, kRowsPerBlock, smem, ctx->stream, loss, rows, offset, index, value, label, w, xw, dual, po, pc, hot_w);
    else WH_LAUNCH((spmv_loss_kernel<true, false>), grid, kRowsPerBlock, smem, ctx->stream, loss, rows, offset, index, value, label, w, xw, dual, po, pc, hot_w);
  } else {
    if (hot_w) WH_LAUNCH((spmv_loss_kernel<false, true>), grid, kRowsPerBlock, kTileCap * 4, ctx->stream, loss, rows, offset, index, value, label, w, xw, dual, po, pc, hot_w);
    else WH_LAUNCH((spmv_loss_kernel<false, false>), grid, kRowsPerBlock, kTileCap * 4, ctx->stream, loss, rows, offset, index, value, label, w, xw, dual, po, pc, hot_w);
  }
  return loss_reduce_launch(ctx, grid, po, nullptr, pc, out, fs, nnz);
}

int spmv_times_launch(whContext* ctx, size_t rows, const size_t* offset, const uint32_t* index,
                      const float* value, const float* x, float* y) {
  if (rows == 0) return WH_OK;
  const unsigned grid = (unsigned)div_up(rows, kRowsPerBlock);
  if (value) {
    const size_t smem = kTileCap * 8;
    WH_CUDA(cudaFuncSetAttribute(spmv_times_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   // per device
    WH_LAUNCH(spmv_times_kernel<true>, grid, kRowsPerBlock, smem, ctx->stream, rows, offset, index, value, x, y);
  } else {
    WH_LAUNCH(spmv_times_kernel<false>, grid, kRowsPerBlock, kTileCap * 4, ctx->stream, rows, offset, index, value, x, y);
  }
  return WH_OK;
}

// ---------------------------------------------------------------------- transposed SpMV --

__global__ void fixed_reset_kernel(FixedScale* fs) { fs->maxx_bits = 0; fs->maxv_bits = 0; }

__global__ void absmax_kernel(const float* __restrict__ v, size_t n, unsigned* target) {
  float m = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    m = fmaxf(m, fabsf(v[i]));
  absmax_commit(target, m);
}

// Picks e with  nnz * max|x| * max|val| * 2^e < 2^62  (no overflow however the terms fall).
__global__ void fixed_scale_kernel(FixedScale* fs, size_t nnz, int has_val) {
  float mx = __uint_as_float(fs->maxx_bits);
  float mv = has_val ? __uint_as_float(fs->maxv_bits) : 1.f;
  double m = (double)mx * (double)mv;
  if (has_val == 2) m *= (double)mv;   // terms x * val^2 (DiFacto's (X.*X)' p)
  int e = 0;
  if (m > 0.0 && isfinite(m)) {
    int em; frexp(m, &em);                    // m < 2^em
    int en = 64 - __clzll((unsigned long long)(nnz | 1));  // nnz < 2^en
    e = 61 - em - en;
  }
  e = max(-120, min(120, e));
  fs->scale = ldexpf(1.f, e);
  fs->inv = ldexp(1.0, -e);
}

// learn/base/spmv.h:95-119: y[index[j]] += x[i] * value[j] for every entry of row i.
template <bool HAS_VAL>
__global__ void __launch_bounds__(kRowsPerBlock)
spmv_trans_kernel(size_t rows, const size_t* __restrict__ offset, const uint32_t* __restrict__ index,
                  const float* __restrict__ value, const float* __restrict__ x, long long* acc,
                  const FixedScale* __restrict__ fs, unsigned copies, size_t stride) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ size_t s_off[kRowsPerBlock + 1];
  uint32_t* s_idx = reinterpret_cast<uint32_t*>(smem);
  float* s_val = reinterpret_cast<float*>(smem + sizeof(uint32_t) * kTileCap);
  const bool staged = stage_rows<HAS_VAL>(rows, offset, index, value, s_off, s_idx, s_val);
  const size_t r = (size_t)blockIdx.x * kRowsPerBlock + threadIdx.x;
  if (r >= rows) return;
  const size_t b = s_off[threadIdx.x], e = s_off[threadIdx.x + 1];
  if (b == e) return;
  // One feature can sit in most rows of a minibatch; atomics on one address serialise in L2
  // (~1.3 ns each, measured), so blocks spread over `copies` replicas of the accumulator array.
  acc += (size_t)(blockIdx.x & (copies - 1)) * stride;
  const float xi = x[r];
  const float scale = fs->scale;
  if (staged) {
    const unsigned lb = (unsigned)(b - s_off[0]), le = (unsigned)(e - s_off[0]);
    for (unsigned j = lb; j < le; ++j) fixed_add(acc + s_idx[j], HAS_VAL ? xi * s_val[j] : xi, scale);
  } else {
    for (size_t j = b; j < e; ++j) fixed_add(acc + index[j], HAS_VAL ? xi * value[j] : xi, scale);
  }
}


// Same product, but lanes of a warp that hit the same output merge their terms first. With one
// thread per row, step j of the loop has every lane on its row's j-th entry -- for field-aligned
// data (one id per field per row) the same field, whose popular values then repeat across the
// lanes: one atomic per distinct output per warp step instead of one per lane. Integer sums are
// associative, so the result is bit-identical to the plain kernel's.
template <bool HAS_VAL>
__global__ void __launch_bounds__(kRowsPerBlock)
spmv_trans_agg_kernel(size_t rows, const size_t* __restrict__ offset, const uint32_t* __restrict__ index,
                      const float* __restrict__ value, const float* __restrict__ x, long long* acc,
                      const FixedScale* __restrict__ fs, unsigned copies, size_t stride) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ size_t s_off[kRowsPerBlock + 1];
  uint32_t* s_idx = reinterpret_cast<uint32_t*>(smem);
  float* s_val = reinterpret_cast<float*>(smem + sizeof(uint32_t) * kTileCap);
  const bool staged = stage_rows<HAS_VAL>(rows, offset, index, value, s_off, s_idx, s_val);
  const size_t r = (size_t)blockIdx.x * kRowsPerBlock + threadIdx.x;
  const bool live = r < rows;
  const size_t b = s_off[threadIdx.x], e = live ? s_off[threadIdx.x + 1] : b;
  acc += (size_t)(blockIdx.x & (copies - 1)) * stride;
  const float xi = live ? x[r] : 0.f;
  const float scale = fs->scale;
  const unsigned len = (unsigned)(e - b);
  const unsigned maxlen = __reduce_max_sync(0xffffffffu, len);
  const unsigned lb = (unsigned)(b - s_off[0]);
  const unsigned lane = threadIdx.x & 31;
  for (unsigned j = 0; j < maxlen; ++j) {
    const bool valid = j < len;
    unsigned c = 0;
    float t = 0.f;
    if (valid) {
      c = staged ? s_idx[lb + j] : index[b + j];
      t = HAS_VAL ? xi * (staged ? s_val[lb + j] : value[b + j]) : xi;
    }
    long long q = __float2ll_rn(t * scale);
    const unsigned g = group_equal_keys(c, valid);
    if (g != (1u << lane)) q = group_sum_i64(g, q);
    if (valid && (__ffs(g) - 1) == (int)lane)
      atomicAdd(reinterpret_cast<unsigned long long*>(acc + c), (unsigned long long)q);
  }
}

// y = fixed-point sum back to fp32 (+ p*z for SpMM); clears the accumulator for the next call.
__global__ void fixed_finish_kernel(long long* acc, size_t n, const FixedScale* __restrict__ fs,
                                    float p, const float* __restrict__ z, float* __restrict__ y,
                                    float mul, unsigned copies, size_t stride, const unsigned* d_n, unsigned ystride) {
  if (d_n) n = *d_n;   // the count may live on the device (number of unique keys of a minibatch)
  const double inv = fs->inv;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const long long a = sum_and_clear_replicas(acc, copies, stride, i);
    float s = (float)((double)a * inv) * mul;
    y[i * ystride] = z ? z[i] * p + s : s;
  }
}

unsigned acc_copies() {
  static int c = -1;
  if (c < 0) { const char* e = getenv("WH_ACC_COPIES"); c = e ? atoi(e) : (int)kAccCopies; if (c < 1 || c > (int)kAccCopies || (c & (c - 1))) c = 1; }
  return (unsigned)c;
}

int ensure_acc(whContext* ctx, size_t n, long long** out) {
  n *= kAccCopies;
  if (ctx->acc.cap < n * sizeof(long long)) {
    WH_TRY(ctx->acc.reserve(n * sizeof(long long)));
    WH_CUDA(cudaMemsetAsync(ctx->acc.p, 0, ctx->acc.cap, ctx->stream));
  }
  *out = ctx->acc.as<long long>();
  return WH_OK;
}

int fixed_prepare(whContext* ctx, FixedScale* fs, const float* x, size_t nx, const float* value,
                  size_t nnz, int vpow) {
  const unsigned cap = (unsigned)ctx->num_sms * 8;
  WH_LAUNCH(fixed_reset_kernel, 1, 1, 0, ctx->stream, fs);
  WH_LAUNCH(absmax_kernel, (unsigned)std::min<size_t>(div_up(nx ? nx : 1, 256), cap), 256, 0, ctx->stream, x, nx, &fs->maxx_bits);
  if (value) WH_LAUNCH(absmax_kernel, (unsigned)std::min<size_t>(div_up(nnz ? nnz : 1, 256), cap), 256, 0, ctx->stream, value, nnz, &fs->maxv_bits);
  WH_LAUNCH(fixed_scale_kernel, 1, 1, 0, ctx->stream, fs, nnz, value ? vpow : 0);
  return WH_OK;
}

__global__ void fixed_unit_x_kernel(FixedScale* fs) { fs->maxx_bits = __float_as_uint(1.f); fs->maxv_bits = 0; }
// The same choice for terms dual * value with |dual| <= 1 known up front: only max|value| has to be found.
int fixed_prepare_unit_x(whContext* ctx, FixedScale* fs, const float* value, size_t nnz) {
  const unsigned cap = (unsigned)ctx->num_sms * 8;
  WH_LAUNCH(fixed_unit_x_kernel, 1, 1, 0, ctx->stream, fs);
  WH_LAUNCH(absmax_kernel, (unsigned)std::min<size_t>(div_up(nnz ? nnz : 1, 256), cap), 256, 0, ctx->stream, value, nnz, &fs->maxv_bits);
  WH_LAUNCH(fixed_scale_kernel, 1, 1, 0, ctx->stream, fs, nnz, 1);
  return WH_OK;
}

// Scale selection + accumulation only; the caller converts (fixed_finish or the fused push).
int spmv_trans_accumulate_launch(whContext* ctx, size_t rows, size_t nnz, const size_t* offset,
                                 const uint32_t* index, const float* value, const float* x,
                                 long long* acc, FixedScale* fs, unsigned copies, size_t stride, bool scale_ready) {
  if (!scale_ready) WH_TRY(fixed_prepare(ctx, fs, x, rows, value, nnz, 1));
  const unsigned grid = (unsigned)div_up(rows, kRowsPerBlock);
  static int mode = -1;   // WH_TRANS_MODE=0 (default): one atomic per entry; 1: merged within the warp first
                          // (measured: merging costs more instructions than the atomics it saves)
  if (mode < 0) { const char* e = getenv("WH_TRANS_MODE"); mode = e ? atoi(e) : 0; }
  if (value) {
    const size_t smem = kTileCap * 8;
    WH_CUDA(cudaFuncSetAttribute(spmv_trans_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   // per device
    WH_CUDA(cudaFuncSetAttribute(spmv_trans_agg_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (mode == 0) WH_LAUNCH(spmv_trans_kernel<true>, grid, kRowsPerBlock, smem, ctx->stream, rows, offset, index, value, x, acc, fs, copies, stride);
    else WH_LAUNCH(spmv_trans_agg_kernel<true>, grid, kRowsPerBlock, smem, ctx->stream, rows, offset, index, value, x, acc, fs, copies, stride);
  } else {
    if (mode == 0) WH_LAUNCH(spmv_trans_kernel<false>, grid, kRowsPerBlock, kTileCap * 4, ctx->stream, rows, offset, index, value, x, acc, fs, copies, stride);
    else WH_LAUNCH(spmv_trans_agg_kernel<false>, grid, kRowsPerBlock, kTileCap * 4, ctx->stream, rows, offset, index, value, x, acc, fs, copies, stride);
  }
  return WH_OK;
}

// Accumulator -> fp32 (and clear), for callers that ran spmv_trans_accumulate_launch themselves.
int fixed_finish_launch(whContext* ctx, long long* acc, size_t n, const FixedScale* fs, float* y, float mul,
                        unsigned copies, size_t stride, const unsigned* d_n, unsigned ystride) {
  if (n == 0) return WH_OK;
  const unsigned g2 = (unsigned)std::min<size_t>(div_up(n, 256), (size_t)ctx->num_sms * 8);
  WH_LAUNCH(fixed_finish_kernel, g2, 256, 0, ctx->stream, acc, n, fs, 0.f, (const float*)nullptr, y, mul, copies, stride, d_n, ystride);
  return WH_OK;
}

int spmv_trans_times_launch(whContext* ctx, size_t rows, size_t nnz, const size_t* offset,
                            const uint32_t* index, const float* value, const float* x, float* y,
                            size_t ysize, long long* acc, FixedScale* fs, float mul, unsigned copies) {
  if (ysize == 0) return WH_OK;
  WH_TRY(spmv_trans_accumulate_launch(ctx, rows, nnz, offset, index, value, x, acc, fs, copies, ysize, false));
  const unsigned g2 = (unsigned)std::min<size_t>(div_up(ysize, 256), (size_t)ctx->num_sms * 8);
  WH_LAUNCH(fixed_finish_kernel, g2, 256, 0, ctx->stream, acc, ysize, fs, 0.f, (const float*)nullptr, y, mul, copies, ysize, (const unsigned*)nullptr, 1u);
  return WH_OK;
}

// ------------------------------------------------------- forward + backward with hot keys --
// hot.cuh. One launch does ScalarLoss::Init (Xw), the per-example part of Evaluate, the dual, and X^T dual for
// binary features: the thread that summed row i turns Xw[i] into dual[i] and then walks the same (still staged)
// codes a second time. One block per SM owns a contiguous band of rows, so that the entries of table keys can be
// added up in a 64 KB shared-memory accumulator and leave the SM once per block instead of once per entry:
// lanes of a warp holding the same slot merge first (ballot match + hardware warp reductions), the group's
// leader adds to shared memory, everything else goes to the global fixed-point accumulators as before.
__global__ void hot_w_kernel(HotView hot, const float* __restrict__ w) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= kHotSlots) return;
  const unsigned r = hot.rank[p];
  hot.w[p] = r == kHotNone ? 0.f : __ldg(w + r);
}
// weight of a code: table keys from shared memory, the rest from the pulled array (generic load, one address)
__device__ __forceinline__ float gather_sw(const float* __restrict__ w, const float* s_w, uint32_t code) {
  const float* p = (code & kHotFlag) ? s_w + (code & (kHotSlots - 1)) : w + code;
  return *p;
}
// the code of an entry: staged codes are final; a band too long for the staging buffer is walked in global memory,
// where entries that are not table keys still hold their key
__device__ __forceinline__ uint32_t code_at(const uint32_t* idx, bool staged, const uint2* __restrict__ dir, unsigned pos) {
  uint32_t c = idx[pos];
  if (!staged && !(c & kHotFlag)) {
    const uint2 e = __ldg(dir + (c >> 5));
    c = e.y + __popc(e.x & ((1u << (c & 31)) - 1u));
  }
  return c;
}
// 64-bit add in shared memory out of two native 32-bit atomics. (atomicAdd on a 64-bit shared word compiles to a
// compare-and-swap spin loop, ATOMS.CAST.SPIN, which collapsed under the table's contention: 600 instructions
// per warp step.) The low word's returned old value tells this add whether IT wrapped; the carry rides on the
// high word's add. Every wrap is carried exactly once, so the 64-bit sum is exact whatever the interleaving.
__device__ __forceinline__ void smem_add_i64(long long* slot, long long v) {
  unsigned* p = reinterpret_cast<unsigned*>(slot);
  const unsigned lo = (unsigned)(unsigned long long)v, hi = (unsigned)((unsigned long long)v >> 32);
  const unsigned old = atomicAdd(p, lo);
  const unsigned carry = (old + lo) < old ? 1u : 0u;
  if (hi + carry != 0u) atomicAdd(p + 1, hi + carry);
}

__global__ void __launch_bounds__(1024)
spmv_fused_hot_kernel(int loss, size_t rows, unsigned rows_per_block, unsigned stage_cap,
                      const size_t* __restrict__ offset, const uint32_t* __restrict__ codes,
                      const float* __restrict__ label, const float* __restrict__ w, HotView hot,
                      float* __restrict__ y, float* __restrict__ dual, double* part_objv, unsigned* part_correct,
                      long long* acc, float scale, unsigned copies, size_t stride, const uint2* __restrict__ dir,
                      FixedScale* fs, double fs_inv) {
  // the push that follows reads the fixed-point scale from device memory: host-known here (|dual| <= 1, binary features)
  if (blockIdx.x == 0 && threadIdx.x == 0) { fs->scale = scale; fs->inv = fs_inv; }
  extern __shared__ __align__(16) unsigned char smem[];
  long long* s_acc = reinterpret_cast<long long*>(smem);                       // kHotSlots
  float* s_w = reinterpret_cast<float*>(s_acc + kHotSlots);                    // kHotSlots: the table keys' weights
  size_t* s_off = reinterpret_cast<size_t*>(s_w + kHotSlots);                  // rows_per_block + 1
  uint32_t* s_idx = reinterpret_cast<uint32_t*>(s_off + rows_per_block + 1);   // stage_cap
  __shared__ double s_o[32];
  __shared__ unsigned s_c[32];
  // the table keys' pulled weights: slot -> rank of the key in this minibatch -> w
  for (int i0 = threadIdx.x; i0 < kHotSlots; i0 += blockDim.x * 4) {   // 4 rank -> weight chains in flight
    unsigned rk[4];
    float wv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { const int i = i0 + u * blockDim.x; rk[u] = i < kHotSlots ? __ldg(hot.rank + i) : kHotNone; }
#pragma unroll
    for (int u = 0; u < 4; ++u) wv[u] = rk[u] == kHotNone ? 0.f : __ldg(w + rk[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u) { const int i = i0 + u * blockDim.x; if (i < kHotSlots) { s_acc[i] = 0; s_w[i] = wv[u]; } }
  }
  const size_t r0 = (size_t)blockIdx.x * rows_per_block;
  for (unsigned t = threadIdx.x; t <= rows_per_block; t += blockDim.x) s_off[t] = offset[min(r0 + t, rows)];
  __syncthreads();
  const size_t begin = s_off[0];
  const size_t cnt = s_off[rows_per_block] - begin;
  const bool staged = cnt <= (size_t)stage_cap;
  // Staging is also the rank pass: an entry that is not a table key still holds its transformed key (the mark pass
  // left it there) and becomes prefix + popc from the localizer's directory on its way into shared memory -- no
  // separate launch that reads and rewrites the codes.
  if (staged) {
    for (unsigned i0 = threadIdx.x; i0 < (unsigned)cnt; i0 += blockDim.x * kStageIlp) {
      uint32_t c[kStageIlp];
      uint2 e[kStageIlp];
#pragma unroll
      for (int u = 0; u < kStageIlp; ++u) { const unsigned i = i0 + u * blockDim.x; c[u] = i < (unsigned)cnt ? __ldcs(codes + begin + i) : kHotFlag; }
#pragma unroll
      for (int u = 0; u < kStageIlp; ++u) if (!(c[u] & kHotFlag)) e[u] = __ldg(dir + (c[u] >> 5));
#pragma unroll
      for (int u = 0; u < kStageIlp; ++u) {
        const unsigned i = i0 + u * blockDim.x;
        if (i < (unsigned)cnt) s_idx[i] = (c[u] & kHotFlag) ? c[u] : e[u].y + __popc(e[u].x & ((1u << (c[u] & 31)) - 1u));
      }
    }
  }
  __syncthreads();
  const uint32_t* idx = staged ? s_idx : codes + begin;   // generic pointer: shared or global
  const size_t r = r0 + threadIdx.x;
  const bool live = threadIdx.x < rows_per_block && r < rows;
  const unsigned lb = live ? (unsigned)(s_off[threadIdx.x] - begin) : 0u;
  const unsigned len = live ? (unsigned)(s_off[threadIdx.x + 1] - s_off[threadIdx.x]) : 0u;
  // forward: sequential fp32 in CSR order, four gathers in flight
  double objv = 0;
  unsigned correct = 0;
  float dl = 0.f;
  if (live) {
    float sum = 0.f;
    unsigned j = 0;
    for (; j + 8 <= len; j += 8) {   // eight gathers in flight, the adds stay in CSR order
      float g[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) g[u] = gather_sw(w, s_w, code_at(idx, staged, dir, lb + j + u));
#pragma unroll
      for (int u = 0; u < 8; ++u) sum += g[u];
    }
    for (; j < len; ++j) sum += gather_sw(w, s_w, code_at(idx, staged, dir, lb + j));
    y[r] = sum;
    loss_row(loss, label[r], sum, &dl, &objv, &correct);
    if (dual) dual[r] = dl;
  }
  // backward: every entry of row i contributes dual[i] (binary features), in fixed point
  const long long q = __float2ll_rn(dl * scale);
  long long* gacc = acc + (size_t)(blockIdx.x & (copies - 1)) * stride;
  const unsigned maxlen = __reduce_max_sync(0xffffffffu, len);
  const unsigned lane = threadIdx.x & 31;
  // Skewed walk: at step j lane l is on entry (j + 2l) mod len of its row. With one thread per row and
  // field-aligned rows, the unskewed walk has all 32 lanes on the SAME field at every step, whose popular value
  // then repeats across the lanes -- up to 16 lanes on one accumulator, serialised inside the atomic. Skewed, the
  // lanes of a step sit on 32 different fields (2 is coprime to the usual odd row length; 39 + 2 is odd, so the
  // shared-memory reads of the codes stay conflict-free). Integer sums do not care about the order.
  unsigned pos = len ? (2u * lane) % len : 0u;
  for (unsigned j = 0; j < maxlen; ++j) {
    const bool valid = j < len;
    const uint32_t c = valid ? code_at(idx, staged, dir, lb + pos) : 0u;
    if (++pos >= len) pos = 0u;
    const bool is_hot = valid && (c & kHotFlag);
    if (valid && !is_hot) atomicAdd(reinterpret_cast<unsigned long long*>(gacc + c), (unsigned long long)q);
    // table keys: straight into shared memory. Lanes of a warp on the same slot are serialised by the hardware
    // inside ONE native atomic; merging them in software first (ballot match + masked warp reductions) cost 600
    // instructions per warp step, because reductions over lane masks that differ within the warp run mask by mask.
    if (is_hot) smem_add_i64(s_acc + (c & (kHotSlots - 1)), q);
  }
  // objective / accuracy partial sums of the block
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    objv += __shfl_down_sync(0xffffffffu, objv, o);
    correct += __shfl_down_sync(0xffffffffu, correct, o);
  }
  if (lane == 0) { s_o[threadIdx.x >> 5] = objv; s_c[threadIdx.x >> 5] = correct; }
  __syncthreads();   // also: every shared-memory sum of the block is in
  if (threadIdx.x == 0) {
    double o = 0; unsigned cc = 0;
    for (unsigned k = 0; k < (blockDim.x + 31) / 32; ++k) { o += s_o[k]; cc += s_c[k]; }
    part_objv[blockIdx.x] = o; part_correct[blockIdx.x] = cc;
  }
  // the block's sums of the table keys join the ordinary accumulators at their keys' ranks
  for (int i = threadIdx.x; i < kHotSlots; i += blockDim.x) {
    const long long v = s_acc[i];
    if (v != 0) atomicAdd(reinterpret_cast<unsigned long long*>(gacc + __ldg(hot.rank + i)), (unsigned long long)v);
  }
}

int hot_w_launch(whContext* ctx, const HotView& hot, const float* w) {
  WH_LAUNCH(hot_w_kernel, kHotSlots / 256, 256, 0, ctx->stream, hot, w);
  return WH_OK;
}

// Binary features, codes from localize_direct_hot(..., resolve = false) and its directory, acc all zero on entry. Leaves Xw in xw, the dual in dual (may
// be NULL), the evaluation partials in `out`, X^T dual in acc (fixed point, scale written to fs for the push).
int spmv_fused_hot_launch(whContext* ctx, int loss, size_t rows, size_t nnz, const size_t* offset, const uint32_t* codes,
                          const float* label, const float* w, const HotView& hot, float* xw, float* dual, EvalOut* out,
                          FixedScale* fs, long long* acc, unsigned copies, size_t stride, const void* directory,
                          whContext* reduce_ctx) {
  // one block per SM if the rows allow it, a thread per row. A block needs a whole SM's shared memory, so a few SMs
  // are left to the kernels of the neighbouring minibatches (the AUC cluster of the previous one holds 8 SMs for
  // ~150 us): with a block for every SM the last blocks had to wait for a free SM and the launch took two waves.
  static int spare = -1;
  if (spare < 0) { const char* e = getenv("WH_FUSED_SPARE_SMS"); spare = e ? atoi(e) : 16; }
  const size_t use_sms = (size_t)std::max(1, ctx->num_sms - spare);
  unsigned rpb = (unsigned)align_up(div_up(rows, use_sms), 32);
  rpb = std::max(32u, std::min(rpb, 1024u));
  const size_t budget = 220 * 1024;
  // ... as long as the band's codes fit the staging buffer (average row length; a band that still overflows is
  // walked in global memory): otherwise narrower bands and more blocks
  const size_t avg_len = std::max<size_t>(1, div_up(nnz, rows));
  while (rpb > 32 && (size_t)rpb * avg_len * 4 + (size_t)kHotSlots * 12 + ((size_t)rpb + 1) * 8 > budget) rpb -= 32;
  const unsigned grid = (unsigned)div_up(rows, rpb);
  const size_t fixed = (size_t)kHotSlots * 12 + ((size_t)rpb + 1) * 8;
  const unsigned cap = (unsigned)((budget - fixed) / 4);
  WH_CUDA(cudaFuncSetAttribute(spmv_fused_hot_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
  WH_TRY(ctx->stage[7].reserve((size_t)grid * 16 + 64));
  double* po = ctx->stage[7].as<double>();
  unsigned* pc = reinterpret_cast<unsigned*>(po + grid);
  float sc; double inv;
  unit_terms_scale(nnz, &sc, &inv);
  WH_LAUNCH(spmv_fused_hot_kernel, grid, rpb, budget, ctx->stream, loss, rows, rpb, cap, offset, codes, label, w, hot, xw,
            dual, po, pc, acc, sc, copies, stride, reinterpret_cast<const uint2*>(directory), fs, inv);
  // The one-block fold of the evaluation partials feeds only the progress numbers and the AUC kernel: with a side
  // context it leaves the chain (pull -> forward/backward -> push) and runs where the AUC runs, after this kernel.
  if (reduce_ctx && reduce_ctx != ctx) {
    if (!ctx->side_ev) WH_CUDA(cudaEventCreateWithFlags(&ctx->side_ev, cudaEventDisableTiming));
    WH_CUDA(cudaEventRecord(ctx->side_ev, ctx->stream));
    WH_CUDA(cudaStreamWaitEvent(reduce_ctx->stream, ctx->side_ev, 0));
    return loss_reduce_launch(reduce_ctx, grid, po, nullptr, pc, out, nullptr, nnz);
  }
  return loss_reduce_launch(ctx, grid, po, nullptr, pc, out, nullptr, nnz);
}

// --------------------------------------------------------------------------------- SpMM --

// learn/base/spmm.h:56-81. One thread per (row, column): entries of the row in CSR order,
// column k of X[index[j], :] gathered as a contiguous dim-wide run.
__global__ void __launch_bounds__(256)
spmm_times_kernel(size_t rows, const size_t* __restrict__ offset, const uint32_t* __restrict__ index,
                  const float* __restrict__ value, const float* __restrict__ x, float* __restrict__ y,
                  int dim) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t r = t / dim;
  const int k = (int)(t % dim);
  if (r >= rows) return;
  float acc = 0.f;
  const size_t b = offset[r], e = offset[r + 1];
  for (size_t j = b; j < e; ++j) {
    float g = __ldg(x + (size_t)index[j] * dim + k);
    if (value) g *= value[j];
    acc += g;
  }
  y[r * dim + k] = acc;  // spmm.h:58 zeroes Y first, so empty rows read 0
}

int spmm_times_launch(whContext* ctx, size_t rows, const size_t* offset, const uint32_t* index,
                      const float* value, const float* x, float* y, int dim) {
  if (rows == 0 || dim <= 0) return WH_OK;
  WH_LAUNCH(spmm_times_kernel, (unsigned)div_up(rows * dim, 256), 256, 0, ctx->stream, rows, offset, index, value, x, y, dim);
  return WH_OK;
}

// learn/base/spmm.h:85-123: Y[index[j], :] += X[i, :] * value[j].
__global__ void __launch_bounds__(256)
spmm_trans_kernel(size_t rows, const size_t* __restrict__ offset, const uint32_t* __restrict__ index,
                  const float* __restrict__ value, const float* __restrict__ x, int dim, long long* acc,
                  const FixedScale* __restrict__ fs) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t r = t / dim;
  const int k = (int)(t % dim);
  if (r >= rows) return;
  const float xi = x[r * dim + k];
  const float scale = fs->scale;
  const size_t b = offset[r], e = offset[r + 1];
  for (size_t j = b; j < e; ++j)
    fixed_add(acc + (size_t)index[j] * dim + k, value ? xi * value[j] : xi, scale);
}

int spmm_trans_times_launch(whContext* ctx, size_t rows, size_t nnz, const size_t* offset,
                            const uint32_t* index, const float* value, const float* x, float p,
                            const float* z, float* y, size_t ysize, int dim, long long* acc,
                            FixedScale* fs) {
  if (ysize == 0 || dim <= 0) return WH_OK;
  WH_TRY(fixed_prepare(ctx, fs, x, rows * dim, value, nnz, 1));
  if (rows) WH_LAUNCH(spmm_trans_kernel, (unsigned)div_up(rows * dim, 256), 256, 0, ctx->stream, rows, offset, index, value, x, dim, acc, fs);
  const unsigned g2 = (unsigned)std::min<size_t>(div_up(ysize, 256), (size_t)ctx->num_sms * 8);
  WH_LAUNCH(fixed_finish_kernel, g2, 256, 0, ctx->stream, acc, ysize, fs, p, (z && p != 0.f) ? z : (const float*)nullptr, y, 1.f, 1u, ysize, (const unsigned*)nullptr, 1u);
  return WH_OK;
}

}  // namespace wh

// ------------------------------------------------------------------------------ C ABI --
using namespace wh;

namespace wh {
// Brings a whLocalBlock to the device (stage slots 0..3) and learns its entry count.
int stage_local(whContext* ctx, const whLocalBlock* D, int mem, DevLocal* out) {
  WH_REQUIRE(ctx && D, "NULL argument");
  WH_REQUIRE(mem == WH_HOST || mem == WH_DEVICE, "bad mem %d", mem);
  out->rows = D->size;
  if (D->size == 0) return WH_OK;
  WH_REQUIRE(D->offset != nullptr, "offset is NULL");
  size_t n;
  if (mem == WH_HOST) n = D->offset[D->size];
  else {
    WH_CUDA(cudaMemcpyAsync(ctx->pinned.p, D->offset + D->size, sizeof(size_t), cudaMemcpyDeviceToHost, ctx->stream));
    WH_CUDA(cudaStreamSynchronize(ctx->stream));
    n = ctx->pinned.as<size_t>()[0];
  }
  out->nnz = n;
  WH_TRY(to_device(ctx, ctx->stage[0], D->offset, D->size + 1, mem, &out->offset));
  WH_TRY(to_device(ctx, ctx->stage[1], D->index, n, mem, &out->index));
  WH_TRY(to_device(ctx, ctx->stage[2], D->value, n, mem, &out->value));
  WH_TRY(to_device(ctx, ctx->stage[3], D->label, D->size, mem, &out->label));
  return WH_OK;
}
}  // namespace wh

extern "C" {

int wh_spmv_times(whContext* ctx, const whLocalBlock* D, int mem, const float* x, size_t xsize, float* y) {
  DevLocal d;
  WH_REQUIRE(ctx && D && y, "wh_spmv_times: NULL argument");
  DeviceGuard g(ctx->device);
  WH_TRY(stage_local(ctx, D, mem, &d));
  if (d.rows == 0) return WH_OK;
  const float* dx; float* dy;
  WH_TRY(to_device(ctx, ctx->stage[4], x, xsize, mem, &dx));
  WH_TRY(out_buffer(ctx, ctx->stage[5], y, d.rows, mem, &dy));
  // rows without entries keep the caller's y (spmv.h:79): bring the old contents along
  if (mem == WH_HOST) WH_CUDA(cudaMemcpyAsync(dy, y, d.rows * 4, cudaMemcpyHostToDevice, ctx->stream));
  WH_TRY(spmv_times_launch(ctx, d.rows, d.offset, d.index, d.value, dx, dy));
  WH_TRY(out_commit(ctx, y, dy, d.rows, mem));
  if (mem == WH_HOST) WH_CUDA(cudaStreamSynchronize(ctx->stream));
  return WH_OK;
}

int wh_spmv_trans_times(whContext* ctx, const whLocalBlock* D, int mem, const float* x, float* y, size_t ysize) {
  DevLocal d;
  WH_REQUIRE(ctx && D && y, "wh_spmv_trans_times: NULL argument");
  DeviceGuard g(ctx->device);
  WH_TRY(stage_local(ctx, D, mem, &d));
  const float* dx; float* dy; long long* acc;
  WH_TRY(to_device(ctx, ctx->stage[4], x, d.rows, mem, &dx));
  WH_TRY(out_buffer(ctx, ctx->stage[5], y, ysize, mem, &dy));
  WH_TRY(ensure_acc(ctx, ysize, &acc));
  FixedScale* fs = reinterpret_cast<FixedScale*>(ctx->small.as<char>() + 1024);
  if (d.rows == 0) { if (ysize) WH_CUDA(cudaMemsetAsync(dy, 0, ysize * 4, ctx->stream)); }
  else WH_TRY(spmv_trans_times_launch(ctx, d.rows, d.nnz, d.offset, d.index, d.value, dx, dy, ysize, acc, fs, 1.f, acc_copies()));
  WH_TRY(out_commit(ctx, y, dy, ysize, mem));
  if (mem == WH_HOST) WH_CUDA(cudaStreamSynchronize(ctx->stream));
  return WH_OK;
}

int wh_spmm_times(whContext* ctx, const whLocalBlock* D, int mem, const float* x, size_t xsize, float* y, int dim) {
  DevLocal d;
  WH_REQUIRE(ctx && D && y, "wh_spmm_times: NULL argument");
  WH_REQUIRE(dim > 0, "wh_spmm_times: dim must be positive");
  DeviceGuard g(ctx->device);
  if (xsize == 0) return WH_OK;  // spmm.h:26: if (x.empty()) return
  WH_TRY(stage_local(ctx, D, mem, &d));
  if (d.rows == 0) return WH_OK;
  const float* dx; float* dy;
  WH_TRY(to_device(ctx, ctx->stage[4], x, xsize, mem, &dx));
  WH_TRY(out_buffer(ctx, ctx->stage[5], y, d.rows * dim, mem, &dy));
  WH_TRY(spmm_times_launch(ctx, d.rows, d.offset, d.index, d.value, dx, dy, dim));
  WH_TRY(out_commit(ctx, y, dy, d.rows * dim, mem));
  if (mem == WH_HOST) WH_CUDA(cudaStreamSynchronize(ctx->stream));
  return WH_OK;
}

int wh_spmm_trans_times(whContext* ctx, const whLocalBlock* D, int mem, const float* x, float p,
                        const float* z, float* y, size_t ysize, int dim) {
  DevLocal d;
  WH_REQUIRE(ctx && D && y, "wh_spmm_trans_times: NULL argument");
  WH_REQUIRE(dim > 0, "wh_spmm_trans_times: dim must be positive");
  DeviceGuard g(ctx->device);
  WH_TRY(stage_local(ctx, D, mem, &d));
  if (d.rows == 0) return WH_OK;  // spmm.h:46: if (x.empty()) return
  const float* dx; const float* dz; float* dy; long long* acc;
  WH_TRY(to_device(ctx, ctx->stage[4], x, d.rows * dim, mem, &dx));
  WH_TRY(to_device(ctx, ctx->stage[6], z, ysize, mem, &dz));
  WH_TRY(out_buffer(ctx, ctx->stage[5], y, ysize, mem, &dy));
  WH_TRY(ensure_acc(ctx, ysize, &acc));
  FixedScale* fs = reinterpret_cast<FixedScale*>(ctx->small.as<char>() + 1024);
  WH_TRY(spmm_trans_times_launch(ctx, d.rows, d.nnz, d.offset, d.index, d.value, dx, p, dz, dy, ysize, dim, acc, fs));
  WH_TRY(out_commit(ctx, y, dy, ysize, mem));
  if (mem == WH_HOST) WH_CUDA(cudaStreamSynchronize(ctx->stream));
  return WH_OK;
}

}  // extern "C"

// ------------------------------------------------------------ allocation-free variant --
//
// wh_spmv_trans_times_static: the transposed product for callers that must not allocate.
//
// Who that is: a worker that shares its GPU with minibatch streams it must not stall
// (cudaMalloc/cudaFree synchronise the whole device, so an allocation in one stream's path
// freezes every other stream until it returns), and a process whose device memory has already
// been handed out in full to some other owner (a framework's caching pool, a co-located
// server's model shard), where a fresh cudaMalloc simply fails. The other entry points grow
// their scratch with cudaMalloc on first use or when a larger minibatch arrives; this one
// never does. Its scratch -- the int64 fixed-point accumulators and the scale record -- lives
// in a fixed array inside the library image, so every GPU that loads the library holds those
// 2 MiB until the process exits, and the number of outputs is capped at
// WH_STATIC_SCRATCH_MAX_OUTPUTS.
//
// The array is one per device and shared by every call: two calls that overlap on the same
// GPU (different streams, or different threads) race on it. Serialise them.
#include <mutex>

namespace {

struct StaticScratch {
  long long acc[WH_STATIC_SCRATCH_MAX_OUTPUTS];
  wh::FixedScale fs;
};
// File-scope device array. In host code this name is only a handle for the runtime: the device
// address has to be asked for with cudaGetSymbolAddress (below), never passed to a kernel as is.
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
StaticScratch* g_static_ptr[kStaticMaxDevices] = {};  // device address of the array, per device
int g_static_sms[kStaticMaxDevices] = {};

// Once per device, with that device current: resolve the array's device address.
int static_scratch_setup(int dev, StaticScratch** out, int* sms) {
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (g_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    WH_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    int n = 0;
    WH_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    g_static_sms[dev] = n;
    g_static_ptr[dev] = static_cast<StaticScratch*>(p);
  }
  *out = g_static_ptr[dev];
  *sms = g_static_sms[dev];
  return WH_OK;
}

}  // namespace

extern "C" int wh_spmv_trans_times_static(void* cuda_stream, const whLocalBlock* D, size_t nnz,
                                          const float* x, float* y, size_t ysize) {
  WH_REQUIRE(D && y, "wh_spmv_trans_times_static: NULL argument");
  WH_REQUIRE(ysize <= (size_t)WH_STATIC_SCRATCH_MAX_OUTPUTS,
             "wh_spmv_trans_times_static: %zu outputs exceed the static scratch limit of %d; "
             "use wh_spmv_trans_times, which allocates its scratch",
             ysize, WH_STATIC_SCRATCH_MAX_OUTPUTS);
  WH_REQUIRE(D->size == 0 || (D->offset && D->index && x), "wh_spmv_trans_times_static: NULL array");
  WH_REQUIRE(nnz < (size_t)UINT32_MAX, "wh_spmv_trans_times_static: %zu entries", nnz);
  if (ysize == 0) return WH_OK;
  int dev = 0;
  WH_CUDA(cudaGetDevice(&dev));
  WH_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "wh_spmv_trans_times_static: device %d not supported", dev);
  StaticScratch* sc = nullptr;
  int sms = 0;
  WH_TRY(static_scratch_setup(dev, &sc, &sms));  // first call on this device resolves the address
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  // a launch context on the stack: no buffers, no events, just the stream and the SM count
  whContext lc;
  lc.device = dev;
  lc.num_sms = sms;
  lc.stream = stream;
  lc.own_stream = false;
  // The array still holds whatever the previous call left: the accumulators must start at zero.
  WH_CUDA(cudaMemsetAsync(sc->acc, 0, ysize * sizeof(long long), stream));
  if (D->size == 0) {
    WH_CUDA(cudaMemsetAsync(y, 0, ysize * sizeof(float), stream));
    return WH_OK;
  }
  return wh::spmv_trans_times_launch(&lc, D->size, nnz, D->offset, D->index, D->value, x, y, ysize,
                                     sc->acc, &sc->fs, 1.f, 1u);
}
