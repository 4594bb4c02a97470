// Hand-written stand-in for the protoc output of
// /root/reference/learn/linear/config.proto (protoc is absent here).
// Same getters/has_* and defaults as config.proto:6-134. Test infrastructure only.
#pragma once
#include <string>
namespace dmlc {
namespace linear {
class Config {
 public:
  enum Loss { SQUARE = 1, LOGIT = 2, SQUARE_HINGE = 4 };
  enum Algo { SGD = 1, ADAGRAD = 2, FTRL = 3 };
#define WH_FIELD(type, name, dflt)                                  \
 public:                                                            \
  type name() const { return name##_; }                             \
  bool has_##name() const { return has_##name##_; }                 \
  void set_##name(type v) { name##_ = v; has_##name##_ = true; }    \
 private:                                                           \
  type name##_ = dflt;                                              \
  bool has_##name##_ = false;
  WH_FIELD(std::string, train_data, "")
  WH_FIELD(std::string, val_data, "")
  WH_FIELD(std::string, data_format, "libsvm")
  WH_FIELD(std::string, model_out, "")
  WH_FIELD(std::string, model_in, "")
  WH_FIELD(std::string, predict_out, "")
  WH_FIELD(Loss, loss, LOGIT)
  WH_FIELD(float, lambda_l1, 1)
  WH_FIELD(float, lambda_l2, 0)
  WH_FIELD(Algo, algo, FTRL)
  WH_FIELD(int, minibatch, 1000)
  WH_FIELD(int, max_data_pass, 10)
  WH_FIELD(float, lr_eta, .01f)
  WH_FIELD(int, save_iter, -1)
  WH_FIELD(int, load_iter, -1)
  WH_FIELD(bool, local_data, false)
  WH_FIELD(int, num_parts_per_file, 10)
  WH_FIELD(int, rand_shuffle, 10)
  WH_FIELD(float, neg_sampling, 1.0f)
  WH_FIELD(bool, prob_predict, true)
  WH_FIELD(float, dropout, 0)
  WH_FIELD(float, print_sec, 1)
  WH_FIELD(float, lr_beta, 1)
  WH_FIELD(int, num_threads, 2)
  WH_FIELD(int, max_concurrency, 2)
  WH_FIELD(bool, key_cache, true)
  WH_FIELD(bool, msg_compression, true)
  WH_FIELD(int, fixed_bytes, 0)
#undef WH_FIELD
};
}  // namespace linear
}  // namespace dmlc
