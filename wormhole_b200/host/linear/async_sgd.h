// The app layer of learn/linear/async_sgd.h on one GPU: AsgdServer (the model shard with its
// SGD / AdaGrad / FTRL handle, :71-213) and AsgdWorker::ProcessMinibatch (:229-305), same names and
// conf keys. The scheduler / ps::App plumbing of the reference (roles, dispatch, ZeroMQ) is not
// reproduced: construct a server, hand it to a worker, feed RowBlock<FeaID> minibatches.
#pragma once
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include "base/localizer.h"
#include "config.pb.h"
#include "loss.h"
#include "penalty.h"
#include "progress.h"
#include "ps.h"
#include "base/minibatch_iter.h"

namespace dmlc {
namespace linear {

using FeaID = ps::Key;

/// AsgdServer::CreateServer<Entry, Handle> (learn/linear/async_sgd.h:200-213): lambda_l1 / lambda_l2
/// from the conf, alpha / beta only when lr_eta / lr_beta are PRESENT (the handle defaults .1 and 1
/// otherwise, :49,205-206). Owns the keys [key_begin, key_end) of the transformed key space.
class AsgdServer {
 public:
  explicit AsgdServer(const Config& conf, uint64_t key_begin = 0, uint64_t key_end = 0, uint64_t hash_slots = 1ull << 24)
      : conf_(conf) {
    whServerConfig c;
    c.algo = conf.algo();
    c.lambda_l1 = conf.lambda_l1(); c.lambda_l2 = conf.lambda_l2();
    c.has_lr_eta = conf.has_lr_eta(); c.lr_eta = conf.lr_eta();
    c.has_lr_beta = conf.has_lr_beta(); c.lr_beta = conf.lr_beta();
    c.key_begin = key_begin;
    c.key_end = key_end ? key_end : ps::FLAGS_max_key;
    // hashed key space: dense shard. Default ps::FLAGS_max_key (all ones): 64-bit keys, open-addressing shard
    c.hash_slots = (c.key_end - c.key_begin > (1ull << 32)) ? hash_slots : 0;
    WH_CHECK(wh_server_create(wh::Ctx(), &c, &server_));
  }
  ~AsgdServer() { wh_server_destroy(server_); }
  whServer* server() const { return server_; }
  /// IterServer::SaveModel / LoadModel (learn/solver/iter_solver.h:99-119): the records of w != 0
  void SaveModel(const std::string& path) const { WH_CHECK(wh_server_save(server_, path.c_str())); }
  void LoadModel(const std::string& path) { WH_CHECK(wh_server_load(server_, path.c_str())); }
  /// ISGDHandle::new_w summed so far = |w|_0 (learn/linear/async_sgd.h:35-41)
  int64_t nnz_w() const { int64_t n = 0; WH_CHECK(wh_server_nnz_w(server_, &n)); return n; }
  /// Entry fields of the given keys
  void State(const std::vector<FeaID>& keys, std::vector<float>* w, std::vector<float>* z, std::vector<float>* cg) const {
    if (w) w->resize(keys.size());
    if (z) z->resize(keys.size());
    if (cg) cg->resize(keys.size());
    WH_CHECK(wh_server_get_state(server_, keys.data(), keys.size(), WH_HOST, w ? w->data() : NULL,
                                 z ? z->data() : NULL, cg ? cg->data() : NULL));
  }
 private:
  Config conf_;
  whServer* server_ = nullptr;
};

/// AsgdWorker (learn/linear/async_sgd.h:229-305). ProcessMinibatch keeps the reference's sequence:
/// Localize -> ZPull -> [callback: Init, Evaluate, report, CalcGrad, ZPush]. With the server on the
/// same GPU all of it is one library call; the progress vector comes back through `reporter`.
class AsgdWorker {
 public:
  using Minibatch = RowBlock<FeaID>;
  AsgdWorker(const Config& conf, AsgdServer* server) : conf_(conf) {
    const int depth = conf.max_concurrency() < 1 ? 1 : (conf.max_concurrency() > 8 ? 8 : conf.max_concurrency());
    WH_CHECK(wh_worker_create_ex(wh::Ctx(), server->server(), conf.loss(), ps::FLAGS_max_key, depth, &worker_));
  }
  ~AsgdWorker() { wh_worker_destroy(worker_); }
  /// called once per finished minibatch with {objv, acc, auc, count, new_ex, new_w}
  std::function<void(const Progress&)> reporter;

  void ProcessMinibatch(const Minibatch& mb, bool train = true) {
    whRowBlock rb;
    rb.size = mb.size; rb.offset = mb.offset; rb.label = mb.label; rb.weight = mb.weight;
    rb.index = mb.index; rb.value = mb.value;
    Progress prog;
    WH_CHECK(wh_worker_process_minibatch(worker_, &rb, WH_HOST, train ? 1 : 0, prog.data.data()));
    if (reporter) reporter(prog);
  }
  /// the same for a minibatch that already lives on the device (MinibatchIter::DeviceValue of the GPU reader)
  void ProcessMinibatch(const data::DeviceMinibatch& mb, bool train = true) {
    Progress prog;
    uint64_t ticket = 0;
    WH_CHECK(wh_worker_submit(worker_, &mb.blk, WH_DEVICE, mb.nnz, train ? 1 : 0, &ticket));
    WH_CHECK(wh_worker_wait(worker_, ticket, prog.data.data()));
    if (reporter) reporter(prog);
  }
  /// The two halves of the above, for max_concurrency minibatches in flight (learn/solver/minibatch_solver.h:307-311):
  /// Submit enqueues, Wait collects in submission order and calls `reporter`.
  uint64_t Submit(const data::DeviceMinibatch& mb, bool train = true) {
    uint64_t ticket = 0;
    WH_CHECK(wh_worker_submit(worker_, &mb.blk, WH_DEVICE, mb.nnz, train ? 1 : 0, &ticket));
    return ticket;
  }
  /// a step of a sharded job in which this rank has no minibatch (wh_worker_submit_idle)
  uint64_t SubmitIdle(bool train = true) {
    uint64_t ticket = 0;
    WH_CHECK(wh_worker_submit_idle(worker_, train ? 1 : 0, &ticket));
    return ticket;
  }
  whWorker* handle() const { return worker_; }
  void Wait(uint64_t ticket) {
    Progress prog;
    WH_CHECK(wh_worker_wait(worker_, ticket, prog.data.data()));
    if (reporter) reporter(prog);
  }
  /// ScalarLoss::Predict (learn/linear/loss.h:55-63): one value per row, sigmoid(Xw) if prob_predict else Xw
  void Predict(const data::DeviceMinibatch& mb, bool prob_predict, std::vector<float>* out) {
    out->resize(mb.blk.size);
    WH_CHECK(wh_worker_predict(worker_, &mb.blk, WH_DEVICE, prob_predict ? 1 : 0, out->data()));
  }
  int depth() const { const int c = conf_.max_concurrency(); return c < 1 ? 1 : (c > 8 ? 8 : c); }
 private:
  Config conf_;
  whWorker* worker_ = nullptr;
};

}  // namespace linear
}  // namespace dmlc
