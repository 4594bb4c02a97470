// The reference's linear.dmlc entry (learn/linear/linear.cc + learn/base/arg_parser.h) on one B200:
//   linear_b200 <conf file> [key=value ...] [-max_key N] [-seed N] [-gpus S] [-max_uniq U]
// reads the conf the way ArgParser does (file, then the arguments as overrides), builds scheduler + server + worker in
// this process and runs MinibatchScheduler::Run. -max_key is ps-lite's gflag (default: all ones = un-hashed 64-bit
// keys); -seed seeds rand(), which the reader's shuffle buffer and negative sampling draw from.
// -gpus S (S > 1): one process per GPU, rank g = worker g + the server of key range g (host/solver/sharded_solver.h;
// the role switch of learn/linear/linear.cc:6-31); needs a hashed key space (-max_key). -max_uniq bounds the unique keys
// of one minibatch (the size of the outboxes the GPUs read from each other; default min(max_key, 128 x minibatch)).
#define WH_DEFINE_FLAGS
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "solver/sharded_solver.h"

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s <conf> [key=value ...] [-max_key N] [-seed N] [-gpus S] [-max_uniq U]\n", argv[0]); return 2; }
  int gpus = 1;
  unsigned seed = 0;
  bool seeded = false;
  size_t max_uniq = 0;
  std::string text;
  if (FILE* f = fopen(argv[1], "r")) {
    char buf[4096]; size_t k;
    while ((k = fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, k);
    fclose(f);
  } else { fprintf(stderr, "[FATAL] nothing read from %s\n", argv[1]); return 3; }
  for (int i = 2; i < argc; ++i) {
    if (!strcmp(argv[i], "-max_key") && i + 1 < argc) { ps::FLAGS_max_key = strtoull(argv[++i], NULL, 10); continue; }
    if (!strcmp(argv[i], "-seed") && i + 1 < argc) { seed = (unsigned)atoi(argv[++i]); seeded = true; srand(seed); continue; }
    if (!strcmp(argv[i], "-gpus") && i + 1 < argc) { gpus = atoi(argv[++i]); continue; }
    if (!strcmp(argv[i], "-max_uniq") && i + 1 < argc) { max_uniq = strtoull(argv[++i], NULL, 10); continue; }
    text += "\n";
    text += argv[i];
  }
  dmlc::linear::Config conf;
  std::string err;
  if (!conf.ParseFromString(text, &err)) { fprintf(stderr, "[FATAL] conf: %s\n", err.c_str()); return 3; }
  if (gpus <= 1) {
    dmlc::solver::MinibatchScheduler sched(conf);
    return sched.Run() ? 0 : 1;
  }
  if (gpus > 8) { fprintf(stderr, "[FATAL] -gpus %d: at most 8 (one box)\n", gpus); return 3; }
  if (!max_uniq) max_uniq = (size_t)std::min<uint64_t>(ps::FLAGS_max_key, (uint64_t)128 * (uint64_t)conf.minibatch());
  return dmlc::solver::RunRanks(gpus, [&](dmlc::solver::Peer* peer) {
    if (seeded) srand(seed + (unsigned)peer->rank);     // every worker process its own rand() sequence
    dmlc::solver::ShardedScheduler sched(conf, peer, max_uniq);
    return sched.Run() ? 0 : 1;
  });
}
