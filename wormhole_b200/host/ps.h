// The slice of the ps-lite v1 surface this path uses, bound to the GPU shards of ONE box:
// ps::Key, ps::FLAGS_max_key, ps::SyncOpts{deps,callback,cmd}, ps::KVWorker<float>::{ZPull,ZPush}
// (call sites: learn/linear/async_sgd.h:277-287). Transport, node roles and filters of ps-lite are
// out of scope; a KVWorker here talks to a whServer on the same GPU.
#pragma once
#include <cstdint>
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
  // pulls vals[i] = w(keys[i]); fires the callback when the values are in place
  int ZPull(const std::shared_ptr<std::vector<Key>>& keys, std::vector<V>* vals, const SyncOpts& opts = SyncOpts()) {
    vals->resize(keys->size());
    WH_CHECK(wh_server_pull(server_, keys->data(), keys->size(), WH_HOST, vals->data()));
    if (opts.callback) opts.callback();
    return ts_++;
  }
  // one push message: Handle::Start, Push per key in order, Finish
  int ZPush(const std::shared_ptr<std::vector<Key>>& keys, const std::shared_ptr<std::vector<V>>& vals,
            const SyncOpts& opts = SyncOpts()) {
    WH_CHECK(wh_server_push(server_, keys->data(), keys->size(), WH_HOST, vals->data()));
    if (opts.callback) opts.callback();
    return ts_++;
  }
 private:
  whServer* server_;
  int ts_ = 0;
};
}  // namespace ps
