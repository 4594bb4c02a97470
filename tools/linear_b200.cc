// The reference's linear.dmlc entry (learn/linear/linear.cc + learn/base/arg_parser.h) on one B200:
//   linear_b200 <conf file> [key=value ...] [-max_key N] [-seed N]
// reads the conf the way ArgParser does (file, then the arguments as overrides), builds scheduler + server + worker in
// this process and runs MinibatchScheduler::Run. -max_key is ps-lite's gflag (default: all ones = un-hashed 64-bit
// keys); -seed seeds rand(), which the reader's shuffle buffer and negative sampling draw from.
#define WH_DEFINE_FLAGS
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "solver/minibatch_solver.h"

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s <conf> [key=value ...] [-max_key N] [-seed N]\n", argv[0]); return 2; }
  std::string text;
  if (FILE* f = fopen(argv[1], "r")) {
    char buf[4096]; size_t k;
    while ((k = fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, k);
    fclose(f);
  } else { fprintf(stderr, "[FATAL] nothing read from %s\n", argv[1]); return 3; }
  for (int i = 2; i < argc; ++i) {
    if (!strcmp(argv[i], "-max_key") && i + 1 < argc) { ps::FLAGS_max_key = strtoull(argv[++i], NULL, 10); continue; }
    if (!strcmp(argv[i], "-seed") && i + 1 < argc) { srand((unsigned)atoi(argv[++i])); continue; }
    text += "\n";
    text += argv[i];
  }
  dmlc::linear::Config conf;
  std::string err;
  if (!conf.ParseFromString(text, &err)) { fprintf(stderr, "[FATAL] conf: %s\n", err.c_str()); return 3; }
  dmlc::solver::MinibatchScheduler sched(conf);
  return sched.Run() ? 0 : 1;
}
