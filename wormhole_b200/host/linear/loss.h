// dmlc::linear::{ScalarLoss, BinClassLoss, LogitLoss, SquareHingeLoss, CreateLoss} with the
// interface of learn/linear/loss.h:20-182. Init keeps a view of the data and the weights;
// Evaluate / CalcGrad run Xw, the metrics, the dual and X^T dual on the GPU (wh_linear_loss).
#pragma once
#include <cmath>
#include <ostream>
#include <vector>

#include "base/spmv.h"
#include "config.pb.h"
#include "progress.h"
#include "wh_runtime.h"

namespace dmlc {
namespace linear {

template <typename V>
class ScalarLoss {
  static_assert(sizeof(V) == 4, "fp32 only");
 public:
  ScalarLoss() : init_(false) {}
  virtual ~ScalarLoss() {}
  /// learn/linear/loss.h:32-39. data must outlive this object (the reference keeps a view too).
  void Init(const RowBlock<unsigned>& data, const std::vector<V>& w, int nt) {
    data_ = data; w_ = w; nt_ = nt;
    Xw_.assign(data_.size, 0);
    init_ = true; evaluated_ = false;
  }
  /// learn/linear/loss.h:42-46 and the overrides :81-86, :100-104, :128-139
  virtual void Evaluate(Progress* prog) {
    Require();
    Run(NULL);
    for (int i = 0; i < 6; ++i) prog->data[i] = prog_[i];
  }
  /// learn/linear/loss.h:49, :106-115, :141-156: grad has one entry per weight
  virtual void CalcGrad(std::vector<V>* grad) {
    Require();
    grad->resize(w_.size());
    Run(grad->data());
  }
  /// learn/linear/loss.h:55-63: one prediction per line (text), sigmoid if prob_out
  virtual void Predict(std::ostream* os, bool prob_out) {
    Require();
    if (!evaluated_) Run(NULL);
    for (V p : Xw_) {
      if (prob_out) (*os) << 1.0 / (1.0 + exp(-p)) << "\n"; else (*os) << p << "\n";
    }
  }
  const std::vector<V>& Xw() { Require(); if (!evaluated_) Run(NULL); return Xw_; }

 protected:
  virtual int Type() const = 0;
  void Require() const { if (!init_) { fprintf(stderr, "[FATAL] ScalarLoss: Init not called\n"); abort(); } }
  void Run(V* grad) {
    whLocalBlock b = WhView(data_);
    WH_CHECK(wh_linear_loss(wh::Ctx(), Type(), &b, WH_HOST, w_.data(), w_.size(), prog_, grad, Xw_.data(), NULL));
    evaluated_ = true;
  }
  bool init_, evaluated_ = false;
  RowBlock<unsigned> data_;
  std::vector<V> w_, Xw_;
  double prog_[6];
  int nt_ = 2;
};

template <typename V> class BinClassLoss : public ScalarLoss<V> {};

/// logistic loss, learn/linear/loss.h:89-116
template <typename V>
class LogitLoss : public BinClassLoss<V> {
 protected:
  int Type() const override { return WH_LOSS_LOGIT; }
};

/// squared hinge loss as the reference computes it, learn/linear/loss.h:119-157
template <typename V>
class SquareHingeLoss : public BinClassLoss<V> {
 protected:
  int Type() const override { return WH_LOSS_SQUARE_HINGE; }
};

/// square loss 1/2 (p - y)^2 as learn/linear/config.proto:30-34 declares it. The reference's class is an empty TODO
/// (learn/linear/loss.h:163-166) and its CreateLoss aborts on it (:178-180); this one computes it (no reference to compare).
template <typename V>
class SquareLoss : public ScalarLoss<V> {
 protected:
  int Type() const override { return WH_LOSS_SQUARE; }
};

/// learn/linear/loss.h:172-182
template <typename V>
static ScalarLoss<V>* CreateLoss(Config::Loss loss) {
  switch (loss) {
    case Config::LOGIT: return new LogitLoss<V>();
    case Config::SQUARE: return new SquareLoss<V>();
    case Config::SQUARE_HINGE: return new SquareHingeLoss<V>();
    default: fprintf(stderr, "[FATAL] unknown type: %d\n", (int)loss); abort();
  }
  return NULL;
}

}  // namespace linear
}  // namespace dmlc
