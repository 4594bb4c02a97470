// dmlc::solver::MinibatchScheduler + MinibatchWorker for learn/linear on one GPU, with the reference's control flow
// and console output (learn/solver/minibatch_solver.h:85-192 Run / Iterate / ShowProgress, :284-315 Process;
// learn/linear/async_sgd.h:304-324 the scheduler's progress header and rows; learn/solver/iter_solver.h:113-117,137-153
// model and prediction file names). ps-lite's node roles collapse into this process: the workload pool hands out the
// parts of the matched files in order, the one worker reads them with the GPU reader (base/minibatch_iter.h) and keeps
// max_concurrency minibatches in flight on the whWorker's lanes, server 0 is the whServer on the same GPU.
#pragma once
#include <dirent.h>
#include <sys/stat.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <deque>
#include <regex>
#include <string>
#include <vector>

#include "linear/async_sgd.h"

namespace dmlc {
namespace solver {

enum WorkloadType { TRAIN = 1, VAL = 2, PRED = 3 };   // learn/base/workload.h

// learn/base/match_file.h: the files of dirname(pattern) whose name matches the regex basename(pattern)
inline std::vector<std::string> MatchFile(const std::string& pattern) {
  struct stat st;
  if (stat(pattern.c_str(), &st) == 0 && S_ISREG(st.st_mode)) return {pattern};
  const size_t pos = pattern.find_last_of("/\\");
  const std::string dir = pos == std::string::npos ? "." : pattern.substr(0, pos);
  const std::regex rx(pos == std::string::npos ? pattern : pattern.substr(pos + 1));
  std::vector<std::string> out;
  if (DIR* d = opendir(dir.c_str())) {
    while (dirent* e = readdir(d))
      if (std::regex_search(e->d_name, rx, std::regex_constants::match_continuous)) out.push_back(dir + "/" + e->d_name);
    closedir(d);
  }
  std::sort(out.begin(), out.end());
  return out;
}

class MinibatchScheduler {
 public:
  explicit MinibatchScheduler(const linear::Config& conf, FILE* out = stdout)
      : conf_(conf), out_(out), server_(conf), worker_(conf, &server_) {
    worker_.reporter = [this](const linear::Progress& p) {
      for (int i = 0; i < 6; ++i) acc_[i] += p.data[i];
    };
  }
  linear::AsgdServer* server() { return &server_; }

  /// learn/solver/minibatch_solver.h:85-147
  bool Run() {
    fprintf(out_, "Connected %d servers and %d workers\n", 1, 1);
    start_ = Now();
    const bool is_predict = !conf_.predict_out().empty();
    if (is_predict && conf_.model_in().empty()) { fprintf(stderr, "[FATAL] should provide model_in for predicting\n"); abort(); }
    int cur_iter = 0;
    if (!conf_.model_in().empty()) {
      if (conf_.load_iter() > 0) {
        fprintf(out_, "Loading model from iter = %d\n", conf_.load_iter());
        cur_iter = conf_.load_iter();
      } else {
        fprintf(out_, "Loading the last model\n");
        cur_iter = -1;
      }
      server_.LoadModel(ModelName(conf_.model_in(), cur_iter));
      Iterate(cur_iter, PRED);
      ++cur_iter;
    }
    if (is_predict) { fprintf(out_, "Prediction is finished!\n"); return true; }
    for (; cur_iter < conf_.max_data_pass(); ++cur_iter) {
      if (Iterate(cur_iter, TRAIN) || Iterate(cur_iter, VAL)) { fprintf(out_, "Hit stop critera\n"); break; }
      if (cur_iter == conf_.max_data_pass() - 1) {
        fprintf(out_, "Hit max number of data passes %d\n", conf_.max_data_pass());
        break;
      }
      if (!conf_.model_out().empty() && conf_.save_iter() > 0 && (cur_iter + 1) % conf_.save_iter() == 0) {
        fprintf(out_, "Saving model for iter = %d\n", cur_iter);
        server_.SaveModel(ModelName(conf_.model_out(), cur_iter));
      }
    }
    if (!conf_.model_out().empty()) {
      fprintf(out_, "Saving the final model\n");
      server_.SaveModel(ModelName(conf_.model_out(), -1));
    }
    fprintf(out_, "Training is finished!\n");
    fflush(out_);
    return true;
  }

  /// learn/solver/iter_solver.h:113-117 with rank 0
  static std::string ModelName(const std::string& base, int iter) {
    std::string name = base;
    if (iter >= 0) name += "_iter-" + std::to_string(iter);
    return name + "_part-0";
  }

 protected:
  /// minibatch_solver.h:58 default: never stop early (learn/difacto overrides it)
  virtual bool Stop(const linear::Progress&, bool) { return false; }

 private:
  static double Now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
  }
  // :185-191
  bool ShowProgress(bool is_train) {
    prog_.data.assign(acc_, acc_ + 6);
    std::fill(acc_, acc_ + 6, 0.0);
    const std::string disp = prog_.PrintStr();
    if (disp.empty()) return false;
    fprintf(out_, "%5.0lf  %s\n", Now() - start_, disp.c_str());
    fflush(out_);
    return Stop(prog_, is_train);
  }
  // :149-183, with the dispatch loop replaced by a walk over the parts
  bool Iterate(int iter, WorkloadType type) {
    const bool is_train = type == TRAIN;
    std::string data;
    if (is_train) {
      data = conf_.train_data();
      fprintf(out_, "Training: iter = %d\n", iter);
    } else {
      data = conf_.val_data();
      if (type == PRED) {
        fprintf(out_, "Predicting\n");
      } else {
        fprintf(out_, "Validating: iter = %d\n", iter);
        if (data.empty()) return false;
      }
    }
    fprintf(out_, "  sec %s\n", linear::Progress::HeadStr().c_str());
    fflush(out_);
    bool stop = false;
    double last = Now();
    for (const std::string& file : MatchFile(data)) {
      for (int k = 0; k < conf_.num_parts_per_file() && !stop; ++k) {
        Process(type, file, k, conf_.num_parts_per_file(), [&]() {
          if (is_train && Now() - last >= conf_.print_sec()) {   // print every print_sec seconds while training
            last = Now();
            if (ShowProgress(true)) stop = true;
          }
        });
      }
    }
    return ShowProgress(is_train) || stop;
  }
  // MinibatchWorker::Process, :284-315
  template <typename Tick>
  void Process(WorkloadType type, const std::string& file, int k, int n, Tick tick) {
    const bool train = type == TRAIN;
    const int mb_size = train ? conf_.minibatch() : 10000000;                    // val_mb_size_, :239
    const int shuffle = train ? conf_.minibatch() * conf_.rand_shuffle() : 0;
    const float neg_sp = train ? conf_.neg_sampling() : 1.0f;
    const int max_mb = type == PRED ? 1 : std::min(worker_.depth(), train ? conf_.max_concurrency() : 10);
    data::MinibatchIter<ps::Key> reader(file.c_str(), k, n, conf_.data_format().c_str(), mb_size, shuffle, neg_sp,
                                        (size_t)8 << 20, worker_.depth() + 2);
    reader.BeforeFirst();
    std::deque<uint64_t> fly;
    FILE* pred = nullptr;
    while (reader.Next()) {
      while ((int)fly.size() >= max_mb) { worker_.Wait(fly.front()); fly.pop_front(); tick(); }   // WaitMinibatch(max_mb)
      const data::DeviceMinibatch mb = reader.DeviceValue();
      if (type == PRED) {
        // learn/linear/async_sgd.h:262-265: Evaluate + report as in a VAL pass, then Predict into
        // <predict_out><input file name>_part-<k> (iter_solver.h:137-153)
        worker_.ProcessMinibatch(mb, false);
        if (!conf_.predict_out().empty()) {
          if (!pred) {
            const size_t pos = file.find_last_of("/\\");
            const std::string base = pos == std::string::npos ? file : file.substr(pos + 1);
            pred = fopen((conf_.predict_out() + base + "_part-" + std::to_string(k)).c_str(), "w");
            if (!pred) { fprintf(stderr, "[FATAL] cannot write predictions\n"); abort(); }
          }
          std::vector<float> p;
          worker_.Predict(mb, conf_.prob_predict(), &p);
          for (float v : p) fprintf(pred, "%g\n", v);
        }
      } else {
        fly.push_back(worker_.Submit(mb, train));
      }
    }
    while (!fly.empty()) { worker_.Wait(fly.front()); fly.pop_front(); tick(); }                 // WaitMinibatch(1)
    if (pred) fclose(pred);
  }

  linear::Config conf_;
  FILE* out_;
  linear::AsgdServer server_;
  linear::AsgdWorker worker_;
  linear::Progress prog_;
  double acc_[6] = {0, 0, 0, 0, 0, 0};
  double start_ = 0;
};

}  // namespace solver
}  // namespace dmlc
