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
int auc_launch(whContext* ctx, size_t rows, const float* label, const float* xw, EvalOut* out);
// the multi-launch radix-sort variant (any size); auc_launch uses it beyond 2^20 examples
int auc_launch_multiblock(whContext* ctx, size_t rows, const float* label, const float* xw, EvalOut* out);

// Host-side closing arithmetic of binary_class_evaluation.h:36-38 and :49-50.
struct EvalHost { double auc, auc_lo, auc_hi, acc, objv, logloss; };
EvalHost eval_finish(const EvalOut& e, size_t rows);

}  // namespace wh
