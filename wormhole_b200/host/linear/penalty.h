// dmlc::linear::L1L2<T>: the proximal operator of learn/linear/penalty.h:36-41 as a host scalar, for
// callers that want to reason about one coordinate. The device handles apply the same formula per key.
#pragma once
#include <cstdio>
#include <cstdlib>
namespace dmlc {
namespace linear {
template <typename T>
class L1L2 {
 public:
  void set_lambda1(T v) { if (v < 0) Die("lambda1 >= 0"); l1_ = v; }
  void set_lambda2(T v) { if (v < 0) Die("lambda2 >= 0"); l2_ = v; }
  T lambda1() const { return l1_; }
  T lambda2() const { return l2_; }
  /// argmin_x 0.5*eta*(x - z/eta)^2 + l1*|x| + l2*x^2 ... soft threshold, then scale
  inline T Solve(T z, T eta) const {
    if (!(eta > 0)) Die("eta > 0");
    if (z <= l1_ && z >= -l1_) return 0;
    return (z > 0 ? z - l1_ : z + l1_) / (eta + l2_);
  }
 private:
  static void Die(const char* m) { fprintf(stderr, "[FATAL] Check failed: %s\n", m); abort(); }
  T l1_ = 0, l2_ = 0;
};
}  // namespace linear
}  // namespace dmlc
