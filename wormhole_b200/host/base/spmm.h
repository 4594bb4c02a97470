// dmlc::SpMM with the interface of learn/base/spmm.h:15-52, forwarding to wh_spmm_*.
#pragma once
#include <vector>

#include "base/spmv.h"

namespace dmlc {
class SpMM {
 public:
  static const int kDefaultNT = 2;
  using SpMat = RowBlock<unsigned>;

  /// y = D * x  (learn/base/spmm.h:23-29); dim = y.size() / D.size
  template <typename V>
  static void Times(const SpMat& D, const std::vector<V>& x, std::vector<V>* y, int nt = kDefaultNT) {
    (void)nt;
    if (x.empty()) return;
    int dim = (int)(y->size() / D.size);
    whLocalBlock b = WhView(D);
    WH_CHECK(wh_spmm_times(wh::Ctx(), &b, WH_HOST, x.data(), x.size(), y->data(), dim));
  }
  /// y = D^T * x  (learn/base/spmm.h:34-38)
  template <typename V>
  static void TransTimes(const SpMat& D, const std::vector<V>& x, std::vector<V>* y, int nt = kDefaultNT) {
    TransTimes<V>(D, x, 0, std::vector<V>(), y, nt);
  }
  /// y = D^T * x + p * z  (learn/base/spmm.h:42-52)
  template <typename V>
  static void TransTimes(const SpMat& D, const std::vector<V>& x, V p, const std::vector<V>& z, std::vector<V>* y,
                         int nt = kDefaultNT) {
    (void)nt;
    if (x.empty()) return;
    int dim = (int)(x.size() / D.size);
    const bool use_z = z.size() == y->size() && p != 0;
    // z may alias y (learn/difacto/loss.h:142 passes V.weight as both): stage a copy
    std::vector<V> zc;
    if (use_z) zc = z;
    whLocalBlock b = WhView(D);
    WH_CHECK(wh_spmm_trans_times(wh::Ctx(), &b, WH_HOST, x.data(), use_z ? p : 0, use_z ? zc.data() : NULL,
                                 y->data(), y->size(), dim));
  }
};
}  // namespace dmlc
