// learn/difacto on one GPU behind the reference's names: difacto::Config (the keys AsyncServer / AsyncWorker / Loss read,
// learn/difacto/config.proto:6-167 with one embedding group), AsyncSGD = AsyncServer + AsyncWorker::ProcessMinibatch
// (learn/difacto/async_sgd.h:302-425) forwarding to whDifacto, and AsyncScheduler::Stop (:31-49).
#pragma once
#include <cstdio>
#include <functional>
#include <vector>

#include "dmlc/data.h"
#include "progress.h"
#include "ps.h"
#include "wh_runtime.h"

namespace dmlc {
namespace difacto {

struct Config {
  // w
  float lambda_l1 = 1, lambda_l2 = 0, lr_eta = .01f, lr_beta = 1;
  bool l1_shrk = true;
  // embedding(0); dim 0 = none
  struct Embedding {
    int dim = 0, threshold = 0;
    float lambda_l2 = 0, lr_eta = .01f, lr_beta = 1, init_scale = .01f, dropout = 0, grad_clipping = 0, grad_normalization = 0;
    bool has_lr_eta = false, has_lr_beta = false;
  } embedding;
  // scheduler
  bool early_stop = false;
  float min_objv_decr = 1e-5f, max_objv = 0;
  bool has_max_objv = false;
};

class AsyncSGD {
 public:
  explicit AsyncSGD(const Config& conf, uint64_t seed = 0) {
    whDifactoConfig c;
    c.lambda_l1 = conf.lambda_l1; c.lambda_l2 = conf.lambda_l2; c.lr_eta = conf.lr_eta; c.lr_beta = conf.lr_beta;
    c.l1_shrk = conf.l1_shrk; c.dim = conf.embedding.dim; c.threshold = conf.embedding.threshold;
    c.v_lambda_l2 = conf.embedding.lambda_l2;
    c.v_has_lr_eta = conf.embedding.has_lr_eta; c.v_lr_eta = conf.embedding.lr_eta;
    c.v_has_lr_beta = conf.embedding.has_lr_beta; c.v_lr_beta = conf.embedding.lr_beta;
    c.init_scale = conf.embedding.init_scale; c.grad_clipping = conf.embedding.grad_clipping;
    c.grad_normalization = conf.embedding.grad_normalization; c.dropout = conf.embedding.dropout;
    c.key_begin = 0; c.key_end = ps::FLAGS_max_key; c.max_key = ps::FLAGS_max_key; c.seed = seed;
    WH_CHECK(wh_difacto_create(wh::Ctx(), &c, &d_));
  }
  ~AsyncSGD() { wh_difacto_destroy(d_); }
  /// called once per minibatch with {objv, auc, objv_w, copc, count, new_ex, new_w, new_V}
  std::function<void(const Progress&)> reporter;
  /// the server's next rand() values, to continue a reference run bit-compatibly (async_sgd.h:254)
  void SetInitStream(const std::vector<int>& rand_values) {
    WH_CHECK(wh_difacto_set_init_stream(d_, rand_values.data(), rand_values.size()));
  }
  /// embedding.dropout > 0: the worker's next rand() values (learn/difacto/loss.h:150-154), or shared = true to draw from
  /// the init stream (worker and server in one process)
  void SetDropoutStream(const std::vector<int>& rand_values, bool shared_with_init = false) {
    WH_CHECK(wh_difacto_set_dropout_stream(d_, rand_values.data(), rand_values.size(), shared_with_init ? 1 : 0));
  }
  /// AsyncWorker::ProcessMinibatch (:363-425)
  void ProcessMinibatch(const RowBlock<ps::Key>& mb, bool train, int data_pass) {
    whRowBlock rb;
    rb.size = mb.size; rb.offset = mb.offset; rb.label = mb.label; rb.weight = mb.weight; rb.index = mb.index; rb.value = mb.value;
    Progress prog;
    WH_CHECK(wh_difacto_process_minibatch(d_, &rb, WH_HOST, train ? 1 : 0, data_pass, prog.data.data()));
    if (reporter) reporter(prog);
  }
  whDifacto* handle() const { return d_; }
 private:
  whDifacto* d_ = nullptr;
};

/// AsyncScheduler::Stop (:31-49)
class StopRule {
 public:
  explicit StopRule(const Config& conf) : conf_(conf) {}
  bool operator()(double objv, double new_ex, bool train) {
    const double cur = objv / new_ex;
    if (train) return conf_.has_max_objv && cur > conf_.max_objv;
    const double diff = pre_val_objv_ - cur;
    pre_val_objv_ = cur;
    if (conf_.early_stop && diff < conf_.min_objv_decr) {
      printf("The decrease of validation objective is smaller than the minimal requirement: %g vs %g\n", diff,
             (double)conf_.min_objv_decr);
      return true;
    }
    return false;
  }
 private:
  Config conf_;
  double pre_val_objv_ = 100;
};

}  // namespace difacto
}  // namespace dmlc
