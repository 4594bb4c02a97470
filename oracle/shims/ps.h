// Single-process stand-in for the ps-lite v1 surface the reference's learn/
// headers use (ps-lite is not vendored: /root/reference/Makefile:36-37 clones
// tag v1). TEST INFRASTRUCTURE ONLY, used to build oracle/_ref/.
//
// What is restated here (and therefore "parity unpinned" by any in-tree test of
// the reference) is the server loop that applies a handle to a message:
//   Start(push, ts, cmd, msg); for keys in message order Push/Pull(key, blob, store[key]); Finish()
// with default-constructed entries on first touch, fixed value length 1 for
// ZPush/ZPull and per-key lengths for ZVPush/ZVPull. Call sites it serves:
// learn/linear/async_sgd.h:211-222,277-287, learn/difacto/async_sgd.h:302-342,379-424.
#pragma once
#include <math.h>  // float overloads of exp/log/sqrt at global scope, as learn/linear/loss.h:4 gets them
#include <unistd.h>
#include <array>
#include <condition_variable>
#include <cstdint>
#include <iostream>
#include <mutex>
#include <thread>
#include <cstring>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <vector>
#include "dmlc/io.h"
#include "dmlc/logging.h"

#define DECLARE_uint64(x) extern uint64_t FLAGS_##x

namespace ps {
typedef uint64_t Key;
DECLARE_uint64(max_key);

template <typename T>
struct Blob {
  T* data;
  size_t size;
  Blob() : data(NULL), size(0) {}
  Blob(T* d, size_t s) : data(d), size(s) {}
  template <typename U> Blob(const Blob<U>& o) : data(o.data), size(o.size) {}
  T& operator[](size_t i) const { return data[i]; }
};

struct NodeInfo {
  static int NumServers() { return 1; }
  static int NumWorkers() { return 1; }
  static int MyRank() { return 0; }
  static bool IsWorker() { return true; }
  static bool IsServer() { return true; }
  static bool IsScheduler() { return false; }
};

static const std::string kWorkerGroup = "all_workers";
static const std::string kServerGroup = "all_servers";

class Task {
 public:
  void set_cmd(int c) { cmd_ = c; }
  int cmd() const { return cmd_; }
  void set_msg(const std::string& m) { msg_ = m; }
  const std::string& msg() const { return msg_; }
 private:
  int cmd_ = 0;
  std::string msg_;
};
struct Message { Task task; std::string sender; };

struct FakeManager {
  void AddNodeFailureHandler(std::function<void(const std::string&)>) {}
};
struct FakeSys { FakeManager& manager() { return m_; } FakeManager m_; };

class App {
 public:
  App() {}
  virtual ~App() {}
  static App* Create(int argc, char* argv[]);
  virtual void ProcessRequest(Message*) {}
  virtual void ProcessResponse(Message*) {}
  virtual bool Run() { return true; }
  int Submit(const Task&, const std::string&) { return 0; }
  void Wait(int) {}
  void Reply(Message*, const Task&) {}
 protected:
  FakeSys sys_;
};

// progress channel: worker/server -> scheduler (learn/solver/iter_solver.h:62,92,132)
struct FakeMonitor {
  std::vector<double> last;          // most recent pushed vector
  std::vector<double> sum;           // running sum
  std::vector<std::vector<double>> log;
  static FakeMonitor& Get() { static FakeMonitor m; return m; }
  void Push(const std::vector<double>& v) {
    last = v; log.push_back(v);
    if (sum.size() < v.size()) sum.resize(v.size(), 0);
    for (size_t i = 0; i < v.size(); ++i) sum[i] += v[i];
  }
  void Clear() { last.clear(); sum.clear(); log.clear(); }
};
template <typename T> struct Root {
  void Get(std::vector<T>* out) { *out = FakeMonitor::Get().sum; FakeMonitor::Get().sum.clear(); }
};
template <typename T> struct Slave {
  void Push(const std::vector<T>& v) { FakeMonitor::Get().Push(v); }
};

struct Filter {
  enum Type { FIXING_FLOAT, KEY_CACHING, COMPRESSING, TRUNCATE_FLOAT };
  Filter* set_num_bytes(int) { return this; }
  Filter* set_clear_cache(bool) { return this; }
};
struct SyncOpts {
  std::vector<int> deps;
  std::function<void()> callback;
  int cmd = 0;
  Filter* AddFilter(Filter::Type) { filters.emplace_back(new Filter()); return filters.back().get(); }
  std::vector<std::shared_ptr<Filter>> filters;
};

// ---- server side -----------------------------------------------------------
class KVStore {
 public:
  virtual ~KVStore() {}
  virtual void Load(dmlc::Stream*) {}
  virtual void Save(dmlc::Stream*) {}
  // fixed value length 1 (ZPush / ZPull)
  virtual void HandlePush(const Key* keys, size_t n, const float* vals, int cmd) = 0;
  virtual void HandlePull(const Key* keys, size_t n, float* vals, int cmd) = 0;
  // per-key lengths (ZVPush / ZVPull)
  virtual void HandleVPush(const Key* keys, size_t n, const float* vals, const int* siz, int cmd) = 0;
  virtual void HandleVPull(const Key* keys, size_t n, std::vector<float>* vals, std::vector<int>* siz, int cmd) = 0;
  static KVStore*& Current() { static KVStore* s = NULL; return s; }
};

template <typename V, typename E, typename H>
class KVStoreSparse : public KVStore {
 public:
  explicit KVStoreSparse(const H& h) : handle_(h) {}
  void HandlePush(const Key* keys, size_t n, const float* vals, int cmd) override {
    handle_.Start(true, ts_++, cmd, NULL);
    for (size_t i = 0; i < n; ++i)
      handle_.Push(keys[i], Blob<const V>(vals + i, 1), data_[keys[i]]);
    handle_.Finish();
  }
  void HandlePull(const Key* keys, size_t n, float* vals, int cmd) override {
    handle_.Start(false, ts_++, cmd, NULL);
    for (size_t i = 0; i < n; ++i) {
      Blob<V> send(vals + i, 1);
      handle_.Pull(keys[i], data_[keys[i]], send);
      if (send.data != vals + i) vals[i] = send.data[0];
    }
    handle_.Finish();
  }
  void HandleVPush(const Key* keys, size_t n, const float* vals, const int* siz, int cmd) override {
    handle_.Start(true, ts_++, cmd, NULL);
    size_t p = 0;
    for (size_t i = 0; i < n; ++i) {
      int len = siz ? siz[i] : 1;
      if (len > 0) handle_.Push(keys[i], Blob<const V>(vals + p, len), data_[keys[i]]);
      p += len;
    }
    handle_.Finish();
  }
  void HandleVPull(const Key* keys, size_t n, std::vector<float>* vals, std::vector<int>* siz, int cmd) override {
    handle_.Start(false, ts_++, cmd, NULL);
    vals->clear(); siz->resize(n);
    std::vector<V> tmp(4096);
    for (size_t i = 0; i < n; ++i) {
      Blob<V> send(tmp.data(), tmp.size());
      handle_.Pull(keys[i], data_[keys[i]], send);
      vals->insert(vals->end(), send.data, send.data + send.size);
      (*siz)[i] = static_cast<int>(send.size);
    }
    handle_.Finish();
  }
  std::map<Key, E>& data() { return data_; }
  H& handle() { return handle_; }
 private:
  H handle_;
  std::map<Key, E> data_;
  int ts_ = 0;
};

template <typename V, typename E, typename H>
class OnlineServer {
 public:
  explicit OnlineServer(const H& h, int = 1, int = 1) {
    store_ = new KVStoreSparse<V, E, H>(h);
    KVStore::Current() = store_;
  }
  KVStore* server() { return store_; }
 private:
  KVStoreSparse<V, E, H>* store_;
};

// ---- worker side -----------------------------------------------------------
template <typename V>
class KVWorker {
 public:
  int ZPush(const std::shared_ptr<std::vector<Key>>& keys,
            const std::shared_ptr<std::vector<V>>& vals, const SyncOpts& opts = SyncOpts()) {
    CHECK_EQ(keys->size(), vals->size());
    KVStore::Current()->HandlePush(keys->data(), keys->size(), vals->data(), opts.cmd);
    if (opts.callback) opts.callback();
    return ts_++;
  }
  int ZPull(const std::shared_ptr<std::vector<Key>>& keys, std::vector<V>* vals,
            const SyncOpts& opts = SyncOpts()) {
    vals->resize(keys->size());
    KVStore::Current()->HandlePull(keys->data(), keys->size(), vals->data(), opts.cmd);
    if (opts.callback) opts.callback();
    return ts_++;
  }
  int ZVPush(const std::shared_ptr<std::vector<Key>>& keys,
             const std::shared_ptr<std::vector<V>>& vals,
             const std::shared_ptr<std::vector<int>>& siz, const SyncOpts& opts = SyncOpts()) {
    CHECK_EQ(keys->size(), siz->size());
    KVStore::Current()->HandleVPush(keys->data(), keys->size(), vals->data(), siz->data(), opts.cmd);
    if (opts.callback) opts.callback();
    return ts_++;
  }
  int ZVPull(const std::shared_ptr<std::vector<Key>>& keys, std::vector<V>* vals,
             std::vector<int>* siz, const SyncOpts& opts = SyncOpts()) {
    KVStore::Current()->HandleVPull(keys->data(), keys->size(), vals, siz, opts.cmd);
    if (opts.callback) opts.callback();
    return ts_++;
  }
 private:
  int ts_ = 0;
};

inline int RunSystem(int*, char***) { return 0; }
}  // namespace ps
