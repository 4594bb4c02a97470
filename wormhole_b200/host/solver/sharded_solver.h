// learn/linear on the GPUs of one box, the host in C++: one process per GPU, rank g = worker g + the server of the
// key range [g*max_key/S, (g+1)*max_key/S) -- the role switch of learn/linear/linear.cc:6-31 (scheduler / server /
// worker by node role) becomes "every rank is a server and a worker, rank 0 also prints the scheduler's lines".
// ps-lite's van is replaced by peer-mapped GPU memory (wh_server_ipc_* / wh_worker_ipc_*: pull and push are served by
// the owners over NVLink); the processes themselves only exchange a few bytes through a hub in the parent process:
// the CUDA IPC handles once, one flag per rank and step ("I have a minibatch"), the progress sums once per pass.
//
// What the workload pool does in the reference (hand file parts to whichever worker asks, learn/base/workload_pool.h)
// is a static round-robin here: part i of the (file, k) list goes to rank i % S. The ranks' parts hold different numbers
// of minibatches, and the barriers between the GPUs are collective: a rank that has run out keeps taking part with
// wh_worker_submit_idle until every rank is done with the pass.
#pragma once
#include <sys/socket.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include <cstring>
#include <memory>

#include "solver/minibatch_solver.h"

namespace dmlc {
namespace solver {

inline bool ReadAll(int fd, void* p, size_t n) {
  char* c = static_cast<char*>(p);
  while (n) { const ssize_t k = read(fd, c, n); if (k <= 0) return false; c += k; n -= (size_t)k; }
  return true;
}
inline bool WriteAll(int fd, const void* p, size_t n) {
  const char* c = static_cast<const char*>(p);
  while (n) { const ssize_t k = write(fd, c, n); if (k <= 0) return false; c += k; n -= (size_t)k; }
  return true;
}

/// A rank's end of the hub.
class Peer {
 public:
  Peer(int fd, int rank, int world) : rank(rank), world(world), fd_(fd) {}
  /// every rank gives `bytes` bytes, every rank receives world * bytes in rank order
  void Allgather(const void* mine, size_t bytes, void* all) {
    const uint32_t n = (uint32_t)bytes;
    if (!WriteAll(fd_, &n, 4) || !WriteAll(fd_, mine, bytes) || !ReadAll(fd_, all, bytes * (size_t)world)) {
      fprintf(stderr, "[FATAL] rank %d: the hub is gone (another rank failed)\n", rank);
      abort();
    }
  }
  const int rank, world;
 private:
  int fd_;
};

/// The parent's relay: collects one message from every rank, sends the concatenation to all; ends when the ranks
/// close their ends. Returns false if a rank left while others were still talking.
inline bool RunHub(const std::vector<int>& fds) {
  std::vector<char> all;
  for (;;) {
    uint32_t bytes = 0;
    size_t gone = 0;
    all.clear();
    for (size_t r = 0; r < fds.size(); ++r) {
      uint32_t n = 0;
      if (!ReadAll(fds[r], &n, 4)) { ++gone; continue; }
      if (gone || (r && n != bytes)) return false;
      bytes = n;
      all.resize(all.size() + n);
      if (!ReadAll(fds[r], all.data() + all.size() - n, n)) return false;
    }
    if (gone == fds.size()) return true;
    if (gone) return false;
    for (int fd : fds) if (!WriteAll(fd, all.data(), all.size())) return false;
  }
}

class ShardedScheduler {
 public:
  ShardedScheduler(const linear::Config& conf, Peer* peer, size_t max_uniq, FILE* out = stdout)
      : conf_(conf), peer_(peer), out_(out),
        server_(conf, Bound(peer->rank, peer->world), Bound(peer->rank + 1, peer->world)), worker_(conf, &server_) {
    if (ps::FLAGS_max_key > (1ull << 32)) { fprintf(stderr, "[FATAL] -gpus N needs a hashed key space (-max_key <= 2^32)\n"); abort(); }
    worker_.reporter = [this](const linear::Progress& p) { for (int i = 0; i < 6; ++i) acc_[i] += p.data[i]; };
    // map every shard and every outbox into every rank (CUDA IPC handles through the hub)
    const int S = peer->world;
    std::vector<char> mine(WH_IPC_HANDLE_BYTES), all((size_t)WH_IPC_HANDLE_BYTES * S);
    WH_CHECK(wh_server_ipc_export(server_.server(), mine.data()));
    peer->Allgather(mine.data(), mine.size(), all.data());
    WH_CHECK(wh_server_ipc_attach(server_.server(), peer->rank, S, ps::FLAGS_max_key, all.data()));
    const size_t hb = (size_t)64 * worker_.depth();
    mine.assign(hb, 0); all.assign(hb * S, 0);
    WH_CHECK(wh_worker_ipc_export(worker_.handle(), max_uniq, mine.data()));
    peer->Allgather(mine.data(), hb, all.data());
    WH_CHECK(wh_worker_ipc_attach(worker_.handle(), all.data()));
  }
  linear::AsgdServer* server() { return &server_; }

  /// learn/solver/minibatch_solver.h:85-147; every rank runs it, rank 0 prints
  bool Run() {
    Say("Connected %d servers and %d workers\n", peer_->world, peer_->world);
    start_ = Now();
    if (!conf_.predict_out().empty()) {
      fprintf(stderr, "[FATAL] predict_out: the PRED workload runs on one GPU (linear_b200 without -gpus)\n");
      abort();
    }
    int cur_iter = 0;
    if (!conf_.model_in().empty()) {
      if (conf_.load_iter() > 0) { Say("Loading model from iter = %d\n", conf_.load_iter()); cur_iter = conf_.load_iter(); }
      else { Say("Loading the last model\n"); cur_iter = -1; }
      server_.LoadModel(ModelName(conf_.model_in(), cur_iter));
      ++cur_iter;
    }
    for (; cur_iter < conf_.max_data_pass(); ++cur_iter) {
      if (Iterate(cur_iter, TRAIN) || Iterate(cur_iter, VAL)) { Say("Hit stop critera\n"); break; }
      if (cur_iter == conf_.max_data_pass() - 1) { Say("Hit max number of data passes %d\n", conf_.max_data_pass()); break; }
      if (!conf_.model_out().empty() && conf_.save_iter() > 0 && (cur_iter + 1) % conf_.save_iter() == 0) {
        Say("Saving model for iter = %d\n", cur_iter);
        server_.SaveModel(ModelName(conf_.model_out(), cur_iter));
      }
    }
    if (!conf_.model_out().empty()) {
      Say("Saving the final model\n");
      server_.SaveModel(ModelName(conf_.model_out(), -1));
    }
    Say("Training is finished!\n");
    if (out_) fflush(out_);
    return true;
  }

  /// learn/solver/iter_solver.h:113-117: server r writes <base>[_iter-k]_part-r
  std::string ModelName(const std::string& base, int iter) const {
    std::string name = base;
    if (iter >= 0) name += "_iter-" + std::to_string(iter);
    return name + "_part-" + std::to_string(peer_->rank);
  }

 private:
  static uint64_t Bound(int r, int world) {
    return (uint64_t)(((unsigned __int128)ps::FLAGS_max_key * (unsigned)r) / (unsigned)world);
  }
  static double Now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
  }
  template <typename... A> void Say(const char* fmt, A... a) { if (peer_->rank == 0 && out_) fprintf(out_, fmt, a...); }
  void Say(const char* text) { if (peer_->rank == 0 && out_) fputs(text, out_); }

  // the scheduler's merged row: every rank's sums, added in rank order
  bool ShowProgress() {
    std::vector<double> all((size_t)6 * peer_->world);
    peer_->Allgather(acc_, sizeof(acc_), all.data());
    std::fill(acc_, acc_ + 6, 0.0);
    prog_.data.assign(6, 0.0);
    for (int r = 0; r < peer_->world; ++r) for (int i = 0; i < 6; ++i) prog_.data[i] += all[(size_t)r * 6 + i];
    const std::string disp = prog_.PrintStr();
    if (disp.empty()) return false;
    if (peer_->rank == 0 && out_) { fprintf(out_, "%5.0lf  %s\n", Now() - start_, disp.c_str()); fflush(out_); }
    return false;   // minibatch_solver.h:58: never stop early
  }

  bool Iterate(int iter, WorkloadType type) {
    const bool train = type == TRAIN;
    const std::string data = train ? conf_.train_data() : conf_.val_data();
    if (train) Say("Training: iter = %d\n", iter);
    else { Say("Validating: iter = %d\n", iter); if (data.empty()) return false; }
    Say("  sec %s\n", linear::Progress::HeadStr().c_str());
    if (peer_->rank == 0 && out_) fflush(out_);
    // this rank's parts
    std::vector<std::pair<std::string, int>> parts;
    int i = 0;
    for (const std::string& file : MatchFile(data))
      for (int k = 0; k < conf_.num_parts_per_file(); ++k, ++i)
        if (i % peer_->world == peer_->rank) parts.emplace_back(file, k);
    // (validation keeps the training minibatch size: one outbox holds one minibatch's keys)
    const int mb_size = conf_.minibatch();
    const int shuffle = train ? conf_.minibatch() * conf_.rand_shuffle() : 0;
    const float neg_sp = train ? conf_.neg_sampling() : 1.0f;
    const int max_mb = std::min(worker_.depth(), train ? conf_.max_concurrency() : 10);
    typedef data::MinibatchIter<ps::Key> Reader;
    std::unique_ptr<Reader> reader;
    size_t next_part = 0;
    std::deque<uint64_t> fly;
    auto drain = [&](size_t keep) { while (fly.size() > keep) { worker_.Wait(fly.front()); fly.pop_front(); } };
    auto next = [&]() -> bool {
      for (;;) {
        if (reader && reader->Next()) return true;
        drain(0);               // the reader's device buffers back the minibatches in flight
        reader.reset();
        if (next_part == parts.size()) return false;
        reader.reset(new Reader(parts[next_part].first.c_str(), parts[next_part].second, conf_.num_parts_per_file(),
                                conf_.data_format().c_str(), mb_size, shuffle, neg_sp, (size_t)8 << 20, worker_.depth() + 2));
        reader->BeforeFirst();
        ++next_part;
      }
    };
    std::vector<unsigned char> flags(peer_->world);
    for (;;) {
      const unsigned char have = next() ? 1 : 0;
      peer_->Allgather(&have, 1, flags.data());
      uint64_t mask = 0;
      for (int r = 0; r < peer_->world; ++r) if (flags[r]) mask |= 1ull << r;
      if (!mask) break;
      WH_CHECK(wh_server_set_active_sources(server_.server(), mask));
      drain((size_t)max_mb - 1);
      fly.push_back(have ? worker_.Submit(reader->DeviceValue(), train) : worker_.SubmitIdle(train));
    }
    drain(0);
    return ShowProgress();
  }

  linear::Config conf_;
  Peer* peer_;
  FILE* out_;
  linear::AsgdServer server_;
  linear::AsgdWorker worker_;
  linear::Progress prog_;
  double acc_[6] = {0, 0, 0, 0, 0, 0};
  double start_ = 0;
};

/// linear.dmlc on `gpus` GPUs: forks one process per GPU (before any CUDA call), relays their messages, and returns the
/// job's exit status. `run(rank, peer)` is the rank's body.
template <typename Body>
int RunRanks(int gpus, Body body) {
  std::vector<int> fds;
  std::vector<pid_t> pids;
  for (int r = 0; r < gpus; ++r) {
    int sv[2];
    if (socketpair(AF_UNIX, SOCK_STREAM, 0, sv) != 0) { perror("socketpair"); return 3; }
    const pid_t pid = fork();
    if (pid < 0) { perror("fork"); return 3; }
    if (pid == 0) {
      close(sv[0]);
      for (int fd : fds) close(fd);
      setenv("WH_DEVICE", std::to_string(r).c_str(), 1);
      Peer peer(sv[1], r, gpus);
      const int rc = body(&peer);
      fflush(stdout);
      close(sv[1]);
      _exit(rc);
    }
    close(sv[1]);
    fds.push_back(sv[0]);
    pids.push_back(pid);
  }
  const bool ok = RunHub(fds);
  for (int fd : fds) close(fd);
  int rc = ok ? 0 : 1;
  for (pid_t p : pids) {
    int st = 0;
    waitpid(p, &st, 0);
    if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) rc = 1;
  }
  return rc;
}

}  // namespace solver
}  // namespace dmlc
