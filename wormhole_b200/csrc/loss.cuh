// Piece (3): loss gradient + per-minibatch metrics. Internal launch API (device pointers).
#pragma once
#include "common.cuh"
#include "radix_sort.cuh"

namespace wh {

// Device-side results of one Evaluate (all integer except the objective, which is reduced in a
// fixed order, so every field is bit-reproducible run to run).
struct EvalOut {
  double objv;                     // sum of per-example loss terms
  double logloss;                  // BinClassEval::LogLoss (only filled by wh_bin_eval)
  unsigned long long correct;      // #examples on the right side of threshold 0
  unsigned long long npos;         // #labels > 0
  unsigned long long ngroups;      // #distinct scores
  unsigned long long area_lo;      // #(pos, neg) pairs with score_pos <  score_neg
  unsigned long long ties;         // #(pos, neg) pairs with score_pos == score_neg
};

// dual may be NULL (evaluation only). want_logloss also fills EvalOut::logloss.
int loss_forward_launch(whContext* ctx, int loss, size_t rows, const float* label, const float* xw,
                        float* dual, EvalOut* out, int want_logloss);
int auc_launch(whContext* ctx, size_t rows, const float* label, const float* xw, EvalOut* out, bool few_sms = false);
// the multi-launch radix-sort variant (any size); auc_launch uses it beyond 2^20 examples
int auc_launch_multiblock(whContext* ctx, size_t rows, const float* label, const float* xw, EvalOut* out);

// The counting AUC (loss.cu). auc_scratch_get: the context's histogram block (zeroed at allocation, left all zero by
// every auc_count_launch); auc_hist_ptr: its 65 536 x u64 histogram, for kernels that add their examples themselves
// (auc_hist_add) and then call auc_count_launch(..., hist_filled = true). nparts > 0 also folds the caller's
// per-block objective / accuracy partial sums into `out` (fixed order), which saves the one-block reduce launch.
int auc_scratch_get(whContext* ctx, void** base);
unsigned long long* auc_hist_ptr(void* scratch_base);
int auc_count_launch(whContext* ctx, size_t n, const float* label, const float* xw, EvalOut* out, bool hist_filled,
                     unsigned nparts, const double* part_objv, const unsigned* part_correct);

#ifdef __CUDACC__
// Order-preserving map float -> uint32 (-0 and +0 compare equal, as in the reference's "<").
__device__ __forceinline__ unsigned ordered_bits(float f) {
  if (f == 0.f) f = 0.f;
  unsigned b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
// one example into the AUC histogram: bin = top 16 bits of the ordered score, positives in the high word
__device__ __forceinline__ void auc_hist_add(unsigned long long* hist, float label, float score) {
  atomicAdd(hist + (ordered_bits(score) >> 16), label > 0 ? (1ull << 32) : 1ull);
}
// exp / log in fp64 rounded to fp32: the correctly rounded fp32 result in all but ~1e-9 of cases (what glibc's
// expf / logf deliver on the reference's side)
__device__ __forceinline__ float exp_rn(float a) { return (float)exp((double)a); }
__device__ __forceinline__ float log_rn(float a) { return (float)log((double)a); }
// One example: learn/linear/loss.h:109-113 (logit dual), :145-149 (squared hinge dual, as written: active when
// the margin is > 1), :131-138 and binary_class_evaluation.h:66-74 (objective terms), :40-51 (accuracy).
__device__ __forceinline__ void loss_row(int loss, float lab, float p, float* dual, double* objv, unsigned* correct) {
  const float y = lab > 0 ? 1.f : -1.f;
  if (loss == WH_LOSS_LOGIT) {
    *dual = -y / (1.f + exp_rn(y * p));
    *objv = (double)log_rn(1.f + exp_rn(-y * p));
  } else if (loss == WH_LOSS_SQUARE) {
    // learn/linear/config.proto:30-34 declares SQUARE = 1 as 1/2 (p - y)^2; the reference never implemented it
    // (learn/linear/loss.h:163-166 is an empty TODO, CreateLoss :178-180 aborts). Here: y = +-1 like the other losses,
    // dual = d/dp = p - y, objective = 1/2 (p - y)^2. No reference to compare with; tests hold it to the closed form.
    const float d = p - y;
    *dual = d;
    *objv = 0.5 * (double)(d * d);
  } else {
    *dual = y * ((y * p > 1.0f) ? 1.f : 0.f);
    const float t = fmaxf(1.f - y * p, 0.f);
    *objv = (double)(t * t);
  }
  *correct = ((lab > 0 && p > 0.f) || (lab <= 0 && p <= 0.f)) ? 1u : 0u;
}
#endif

// Xw, the dual and the objective / accuracy partial sums in ONE pass over the rows (SpMV::Times with the loss as
// its epilogue), then the one-block reduction. With fs != NULL the reduction also writes the fixed-point scale for
// the transposed product that follows, for |dual| <= 1 and binary features (no max-reduction kernels needed).
int spmv_loss_launch(whContext* ctx, int loss, size_t rows, size_t nnz, const size_t* offset, const uint32_t* index,
                     const float* value, const float* label, const float* w, float* xw, float* dual, EvalOut* out,
                     struct FixedScale* fs, const float* hot_w = nullptr);   // hot_w: index holds codes (hot.cuh)
int loss_reduce_launch(whContext* ctx, unsigned nparts, const double* po, const double* pl, const unsigned* pc,
                       EvalOut* out, struct FixedScale* fs, size_t nnz);

// Host-side closing arithmetic of binary_class_evaluation.h:36-38 and :49-50.
struct EvalHost { double auc, auc_lo, auc_hi, acc, objv, logloss; };
EvalHost eval_finish(const EvalOut& e, size_t rows);

}  // namespace wh
