// The host-side plumbing of `linear_b200 -gpus S` without any GPU: RunRanks forks S ranks, each Allgathers messages of
// several sizes through the parent's hub and checks that it received every rank's bytes in rank order; one scenario has
// a rank leave early, which must fail the job instead of hanging it. (host/solver/sharded_solver.h)
#define WH_DEFINE_FLAGS
#include <cstdio>
#include <cstring>
#include <vector>

#include "solver/sharded_solver.h"

using dmlc::solver::Peer;
using dmlc::solver::RunRanks;

int main(int argc, char** argv) {
  const int S = argc > 1 ? atoi(argv[1]) : 3;
  const bool quitter = argc > 2 && !strcmp(argv[2], "quit");
  const int rc = RunRanks(S, [&](Peer* peer) -> int {
    for (int round = 0; round < 50; ++round) {
      if (quitter && peer->rank == 1 && round == 7) return 0;       // leaves while the others still talk
      const size_t bytes = 1 + (size_t)(round * 37) % 300;
      std::vector<unsigned char> mine(bytes), all(bytes * (size_t)S);
      for (size_t i = 0; i < bytes; ++i) mine[i] = (unsigned char)(peer->rank * 31 + round + i);
      peer->Allgather(mine.data(), bytes, all.data());
      for (int r = 0; r < S; ++r)
        for (size_t i = 0; i < bytes; ++i)
          if (all[(size_t)r * bytes + i] != (unsigned char)(r * 31 + round + i)) { fprintf(stderr, "rank %d: bad byte\n", peer->rank); return 1; }
    }
    // the step flags of ShardedScheduler::Iterate: one byte per rank
    unsigned char have = peer->rank % 2, flags[8];
    peer->Allgather(&have, 1, flags);
    for (int r = 0; r < S; ++r) if (flags[r] != r % 2) return 1;
    return 0;
  });
  printf("hub_test %s\n", rc == 0 ? "ok" : "failed");
  return rc;
}
