// dmlc::linear::Config without protobuf: the getters / has_*() the worker, server and scheduler
// code calls, the defaults of learn/linear/config.proto:6-134, and a parser for the same text
// format the reference reads (conf file lines "key = value", '#' comments, CLI overrides
// "key=value"; learn/base/arg_parser.h:18-60).
#pragma once
#include <cstdlib>
#include <sstream>
#include <string>

namespace dmlc {
namespace linear {

class Config {
 public:
  enum Loss { SQUARE = 1, LOGIT = 2, SQUARE_HINGE = 4 };
  enum Algo { SGD = 1, ADAGRAD = 2, FTRL = 3 };

#define WH_CONF_FIELD(type, name, dflt)                          \
 public:                                                         \
  type name() const { return name##_; }                          \
  bool has_##name() const { return has_##name##_; }              \
  void set_##name(type v) { name##_ = v; has_##name##_ = true; } \
 private:                                                        \
  type name##_ = dflt;                                           \
  bool has_##name##_ = false;
  WH_CONF_FIELD(std::string, train_data, "")
  WH_CONF_FIELD(std::string, val_data, "")
  WH_CONF_FIELD(std::string, data_format, "libsvm")
  WH_CONF_FIELD(std::string, model_out, "")
  WH_CONF_FIELD(std::string, model_in, "")
  WH_CONF_FIELD(std::string, predict_out, "")
  WH_CONF_FIELD(Loss, loss, LOGIT)
  WH_CONF_FIELD(float, lambda_l1, 1)
  WH_CONF_FIELD(float, lambda_l2, 0)
  WH_CONF_FIELD(Algo, algo, FTRL)
  WH_CONF_FIELD(int, minibatch, 1000)
  WH_CONF_FIELD(int, max_data_pass, 10)
  WH_CONF_FIELD(float, lr_eta, .01f)
  WH_CONF_FIELD(int, save_iter, -1)
  WH_CONF_FIELD(int, load_iter, -1)
  WH_CONF_FIELD(bool, local_data, false)
  WH_CONF_FIELD(int, num_parts_per_file, 10)
  WH_CONF_FIELD(int, rand_shuffle, 10)
  WH_CONF_FIELD(float, neg_sampling, 1.0f)
  WH_CONF_FIELD(bool, prob_predict, true)
  WH_CONF_FIELD(float, dropout, 0)
  WH_CONF_FIELD(float, print_sec, 1)
  WH_CONF_FIELD(float, lr_beta, 1)
  WH_CONF_FIELD(int, num_threads, 2)
  WH_CONF_FIELD(int, max_concurrency, 2)
  WH_CONF_FIELD(bool, key_cache, true)
  WH_CONF_FIELD(bool, msg_compression, true)
  WH_CONF_FIELD(int, fixed_bytes, 0)
#undef WH_CONF_FIELD

 public:
  /// Parses conf text (one "key = value" or "key: value" per line). Returns false and sets *err on
  /// an unknown key or a bad enum, where protobuf's TextFormat would fail too.
  bool ParseFromString(const std::string& text, std::string* err = nullptr) {
    std::istringstream in(text);
    std::string line;
    while (std::getline(in, line)) {
      size_t h = line.find('#');
      if (h != std::string::npos) line.erase(h);
      size_t eq = line.find_first_of("=:");
      if (eq == std::string::npos) { if (Trim(line).empty()) continue; return Fail(err, "no '=' in: " + line); }
      std::string k = Trim(line.substr(0, eq)), v = Trim(line.substr(eq + 1));
      if (v.size() >= 2 && v.front() == '"' && v.back() == '"') v = v.substr(1, v.size() - 2);
      if (!Set(k, v)) return Fail(err, "unknown key or bad value: " + k + " = " + v);
    }
    return true;
  }

 private:
  static std::string Trim(const std::string& s) {
    size_t b = s.find_first_not_of(" \t\r\n"), e = s.find_last_not_of(" \t\r\n");
    return b == std::string::npos ? "" : s.substr(b, e - b + 1);
  }
  static bool Fail(std::string* err, const std::string& m) { if (err) *err = m; return false; }
  static bool B(const std::string& v) { return v == "true" || v == "1"; }
  bool Set(const std::string& k, const std::string& v) {
#define S_STR(n) if (k == #n) { set_##n(v); return true; }
#define S_INT(n) if (k == #n) { set_##n(atoi(v.c_str())); return true; }
#define S_FLT(n) if (k == #n) { set_##n((float)atof(v.c_str())); return true; }
#define S_BOOL(n) if (k == #n) { set_##n(B(v)); return true; }
    S_STR(train_data) S_STR(val_data) S_STR(data_format) S_STR(model_out) S_STR(model_in) S_STR(predict_out)
    S_FLT(lambda_l1) S_FLT(lambda_l2) S_INT(minibatch) S_INT(max_data_pass) S_FLT(lr_eta) S_INT(save_iter)
    S_INT(load_iter) S_BOOL(local_data) S_INT(num_parts_per_file) S_INT(rand_shuffle) S_FLT(neg_sampling)
    S_BOOL(prob_predict) S_FLT(dropout) S_FLT(print_sec) S_FLT(lr_beta) S_INT(num_threads) S_INT(max_concurrency)
    S_BOOL(key_cache) S_BOOL(msg_compression) S_INT(fixed_bytes)
#undef S_STR
#undef S_INT
#undef S_FLT
#undef S_BOOL
    if (k == "loss") {
      if (v == "SQUARE") set_loss(SQUARE); else if (v == "LOGIT") set_loss(LOGIT);
      else if (v == "SQUARE_HINGE") set_loss(SQUARE_HINGE); else return false;
      return true;
    }
    if (k == "algo") {
      if (v == "SGD") set_algo(SGD); else if (v == "ADAGRAD") set_algo(ADAGRAD);
      else if (v == "FTRL") set_algo(FTRL); else return false;
      return true;
    }
    return false;
  }
};

}  // namespace linear
}  // namespace dmlc
