// dmlc::SpMV with the interface of learn/base/spmv.h:46-120, forwarding to wh_spmv_*.
#pragma once
#include <vector>

#include "dmlc/data.h"
#include "wh_runtime.h"

namespace dmlc {

/// learn/base/spmv.h:10-41
struct Range {
  Range(size_t b, size_t e) : begin(b), end(e) {}
  Range() : begin(0), end(0) {}
  inline Range Segment(size_t idx, size_t nparts) const {
    double itv = static_cast<double>(end - begin) / static_cast<double>(nparts);
    size_t b = static_cast<size_t>(begin + itv * idx);
    size_t e = (idx == nparts - 1) ? end : static_cast<size_t>(begin + itv * (idx + 1));
    return Range(b, e);
  }
  inline bool Has(size_t i) const { return begin <= i && i < end; }
  size_t begin, end;
};

inline whLocalBlock WhView(const RowBlock<unsigned>& D) {
  whLocalBlock b;
  b.size = D.size; b.offset = D.offset; b.label = D.label; b.index = D.index; b.value = D.value;
  return b;
}

class SpMV {
 public:
  static const int kDefaultNT = 2;
  using SpMat = RowBlock<unsigned>;

  /// y = D * x  (learn/base/spmv.h:54-59)
  template <typename V>
  static void Times(const SpMat& D, const std::vector<V>& x, std::vector<V>* y, int nthreads = kDefaultNT) {
    if (y->size() != D.size) { fprintf(stderr, "[FATAL] SpMV::Times: y.size() != D.size\n"); abort(); }
    Times<V>(D, x.data(), y->data(), nthreads, x.size());
  }
  /// y = D^T * x  (learn/base/spmv.h:63-68)
  template <typename V>
  static void TransTimes(const SpMat& D, const std::vector<V>& x, std::vector<V>* y, int nthreads = kDefaultNT) {
    if (x.size() != D.size) { fprintf(stderr, "[FATAL] SpMV::TransTimes: x.size() != D.size\n"); abort(); }
    TransTimes<V>(D, x.data(), y->data(), y->size(), nthreads);
  }
  /// learn/base/spmv.h:72-91. xsize: entries of x (the raw-pointer form of the reference does not
  /// carry it; pass max index + 1, or use the vector form)
  template <typename V>
  static void Times(const SpMat& D, const V* const x, V* y, int nthreads = kDefaultNT, size_t xsize = 0) {
    static_assert(sizeof(V) == 4, "fp32 only");
    (void)nthreads;
    whLocalBlock b = WhView(D);
    WH_CHECK(wh_spmv_times(wh::Ctx(), &b, WH_HOST, x, xsize ? xsize : MaxIndex(D) + 1, y));
  }
  /// learn/base/spmv.h:95-119
  template <typename V>
  static void TransTimes(const SpMat& D, const V* const x, V* y, size_t y_size, int nthreads = kDefaultNT) {
    static_assert(sizeof(V) == 4, "fp32 only");
    (void)nthreads;
    whLocalBlock b = WhView(D);
    WH_CHECK(wh_spmv_trans_times(wh::Ctx(), &b, WH_HOST, x, y, y_size));
  }

 private:
  static size_t MaxIndex(const SpMat& D) {
    size_t m = 0;
    for (size_t j = 0; D.size && j < D.offset[D.size]; ++j) if (D.index[j] > m) m = D.index[j];
    return m;
  }
};
}  // namespace dmlc
