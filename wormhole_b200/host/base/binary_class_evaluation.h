// dmlc::BinClassEval<V> with the interface of learn/base/binary_class_evaluation.h:8-92.
#pragma once
#include "wh_runtime.h"
namespace dmlc {
template <typename V>
class BinClassEval {
  static_assert(sizeof(V) == 4, "fp32 only");
 public:
  BinClassEval(const V* const label, const V* const predict, size_t n, int num_threads = 2)
      : label_(label), predict_(predict), size_(n) { (void)num_threads; }
  V AUC() { Run(); return out_[0]; }
  V Accuracy(V threshold) {
    if (threshold != 0) { fprintf(stderr, "[FATAL] BinClassEval::Accuracy: only threshold 0 is used on this path\n"); abort(); }
    Run(); return out_[1];
  }
  V LogLoss() { Run(); return out_[2]; }
  V LogitObjv() { Run(); return out_[3]; }
  // bounds over every order std::sort may leave equal scores in (no counterpart in the reference)
  V AUCLowerBound() { Run(); return out_[4]; }
  V AUCUpperBound() { Run(); return out_[5]; }
 private:
  void Run() {
    if (done_) return;
    WH_CHECK(wh_bin_eval(wh::Ctx(), label_, predict_, size_, WH_HOST, out_));
    done_ = true;
  }
  V const* label_; V const* predict_; size_t size_;
  float out_[6]; bool done_ = false;
};
}  // namespace dmlc
