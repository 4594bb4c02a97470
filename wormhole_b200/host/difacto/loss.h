// difacto::Loss<float>, the interface of learn/difacto/loss.h:24-158, forwarding to wh_difacto_loss: constructed from
// the localized minibatch and the ZVPull payload (model, model_siz), Evaluate fills the progress vector, CalcGrad
// overwrites the payload with the gradients in the same layout (what the worker then hands to ZVPush).
#pragma once
#include <cstdlib>
#include <vector>

#include "async_sgd.h"
#include "data/row_block.h"

namespace dmlc {
namespace difacto {

template <typename T>
class Loss {
  static_assert(sizeof(T) == 4, "the GPU path computes in float");
 public:
  Loss(const RowBlock<unsigned>& data, const std::vector<T>& model, const std::vector<int>& model_siz, const Config& conf)
      : data_(data), model_(model), siz_(model_siz), conf_(conf) {}

  /// learn/difacto/loss.h:82-96: objv, auc, objv_w (the linear part alone), count
  void Evaluate(Progress* prog) { Run(prog->data.data(), NULL); }

  /// learn/difacto/loss.h:104-158; *grad takes the place of the model, length and layout unchanged. With
  /// embedding.dropout > 0 the decisions come from THIS process's rand(), one call per embedding-gradient float in key
  /// order, exactly as :150-154 draws them (the host is the worker process).
  void CalcGrad(std::vector<T>* grad) {
    double prog[8];
    std::vector<T> g(model_.size());
    Run(prog, g.empty() ? NULL : g.data());
    grad->swap(g);
  }

 private:
  void Run(double* prog, T* grad) {
    whLocalBlock lb;
    lb.size = data_.size; lb.offset = data_.offset; lb.label = data_.label; lb.index = data_.index; lb.value = data_.value;
    const int dim = conf_.embedding.dim;
    std::vector<int> draws;
    if (grad && dim > 0 && conf_.embedding.dropout > 0) {
      size_t need = 0;
      for (int s : siz_) if (s == dim + 1) need += (size_t)dim;
      draws.resize(need);
      for (size_t i = 0; i < need; ++i) draws[i] = rand();
    }
    WH_CHECK(wh_difacto_loss_dropout(wh::Ctx(), &lb, model_.data(), model_.size(), siz_.data(), siz_.size(), dim,
                                     conf_.embedding.grad_clipping, conf_.embedding.grad_normalization,
                                     draws.empty() ? 0.f : conf_.embedding.dropout, draws.data(), draws.size(), prog, grad));
  }
  RowBlock<unsigned> data_;
  const std::vector<T>& model_;
  const std::vector<int>& siz_;
  Config conf_;
};

}  // namespace difacto
}  // namespace dmlc
