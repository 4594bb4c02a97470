// The slice of the ps-lite v1 surface this path uses, bound to the GPU shards of ONE box:
// ps::Key, ps::FLAGS_max_key, ps::SyncOpts{deps,callback,cmd}, ps::KVWorker<float>::{ZPull,ZPush,ZVPull,ZVPush,Wait}
// (call sites: learn/linear/async_sgd.h:277-287, learn/difacto/async_sgd.h:374-381,408-411,424). Transport, node
// roles and filters of ps-lite are out of scope; a KVWorker here talks to a whServer (linear) or a whDifacto shard
// (factorization machine) on the same GPU.
//
// Timestamps: every call returns the timestamp of its message, as ps-lite does. The calls are synchronous -- a
// message is applied, and its callback has run, before the call returns -- so a dependency listed in SyncOpts::deps is
// already satisfied when it can be named at all; the facade checks that every dep IS a timestamp it has issued (a dep
// from the future would deadlock ps-lite) and Wait(ts) returns at once for those.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <memory>
#include <vector>

#include "wh_runtime.h"

namespace ps {
typedef uint64_t Key;
// ps-lite's gflag; all-ones = "no hashing" (learn/base/localizer.h:10,108-121). One definition per program.
extern uint64_t FLAGS_max_key;
#ifdef WH_DEFINE_FLAGS
uint64_t FLAGS_max_key = ~0ULL;
#endif

struct SyncOpts {
  std::vector<int> deps;
  std::function<void()> callback;
  int cmd = 0;
};

template <typename V>
class KVWorker {
 public:
  explicit KVWorker(whServer* server = nullptr) : server_(server) {}
  void set_server(whServer* s) { server_ = s; }
  // the DiFacto shard behind ZVPull / ZVPush and the count push (cmd = kPushFeaCnt = 1, learn/difacto/async_sgd.h:59)
  void set_difacto(whDifacto* d, int dim) { difacto_ = d; dim_ = dim; }
  static const int kPushFeaCnt = 1;
  // pulls vals[i] = w(keys[i]); fires the callback when the values are in place
  int ZPull(const std::shared_ptr<std::vector<Key>>& keys, std::vector<V>* vals, const SyncOpts& opts = SyncOpts()) {
    CheckDeps(opts);
    vals->resize(keys->size());
    WH_CHECK(wh_server_pull(server_, keys->data(), keys->size(), WH_HOST, vals->data()));
    if (opts.callback) opts.callback();
    return ts_++;
  }
  // one push message: Handle::Start, Push per key in order, Finish
  int ZPush(const std::shared_ptr<std::vector<Key>>& keys, const std::shared_ptr<std::vector<V>>& vals,
            const SyncOpts& opts = SyncOpts()) {
    CheckDeps(opts);
    if (difacto_ && opts.cmd == kPushFeaCnt) {
      // the feature counts of the first pass (learn/difacto/async_sgd.h:374-379): one float per key
      Require(vals->size() == keys->size(), "ZPush(kPushFeaCnt): one count per key");
      WH_CHECK(wh_difacto_zpush_count(difacto_, keys->data(), vals->data(), keys->size()));
    } else {
      WH_CHECK(wh_server_push(server_, keys->data(), keys->size(), WH_HOST, vals->data()));
    }
    if (opts.callback) opts.callback();
    return ts_++;
  }
  // variable-length values: sizes[i] floats for keys[i], concatenated in vals (learn/difacto/async_sgd.h:424)
  int ZVPull(const std::shared_ptr<std::vector<Key>>& keys, std::vector<V>* vals, std::vector<int>* sizes,
             const SyncOpts& opts = SyncOpts()) {
    CheckDeps(opts);
    Require(difacto_, "ZVPull: no DiFacto shard set (set_difacto)");
    const size_t n = keys->size();
    sizes->resize(n);
    vals->resize(n * (size_t)(dim_ + 1));
    size_t len = 0;
    WH_CHECK(wh_difacto_zvpull(difacto_, keys->data(), n, vals->data(), sizes->data(), &len));
    vals->resize(len);
    if (opts.callback) opts.callback();
    return ts_++;
  }
  // ... and the gradients back in the same layout (learn/difacto/async_sgd.h:408-411)
  int ZVPush(const std::shared_ptr<std::vector<Key>>& keys, const std::shared_ptr<std::vector<V>>& vals,
             const std::shared_ptr<std::vector<int>>& sizes, const SyncOpts& opts = SyncOpts()) {
    CheckDeps(opts);
    Require(difacto_, "ZVPush: no DiFacto shard set (set_difacto)");
    Require(sizes->size() == keys->size(), "ZVPush: one length per key");
    WH_CHECK(wh_difacto_zvpush(difacto_, keys->data(), keys->size(), vals->data(), sizes->data(), vals->size()));
    if (opts.callback) opts.callback();
    return ts_++;
  }
  // blocks until message `timestamp` is done: it is, if it was ever issued
  void Wait(int timestamp) { Require(timestamp >= 0 && timestamp < ts_, "Wait: a timestamp this worker has not issued"); }

 private:
  void CheckDeps(const SyncOpts& opts) {
    for (int t : opts.deps) Require(t >= 0 && t < ts_, "SyncOpts.deps names a timestamp this worker has not issued");
  }
  // glog CHECK behaviour, like every other failure of the facade (wh_runtime.h)
  static void Require(bool ok, const char* what) { if (!ok) { fprintf(stderr, "[FATAL] ps::KVWorker: %s\n", what); abort(); } }
  whServer* server_;
  whDifacto* difacto_ = nullptr;
  int dim_ = 0;
  int ts_ = 0;
};
}  // namespace ps
