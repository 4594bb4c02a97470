// Prints, for every part of a text file, the chunks dmlc::data::detail::TextSplit serves (size and first / last byte),
// without touching the GPU: tests/test_solver_host.py checks that parts and chunks tile the file at line starts.
#define WH_DEFINE_FLAGS
#include <cstdio>
#include <cstdlib>

#include "base/minibatch_iter.h"

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  const unsigned nparts = (unsigned)atoi(argv[2]);
  const size_t chunk = (size_t)atoll(argv[3]);
  for (unsigned k = 0; k < nparts; ++k) {
    dmlc::data::detail::TextSplit split(argv[1], k, nparts, chunk);
    std::vector<char> buf;
    while (split.NextChunk(&buf)) printf("%u %zu %d %d\n", k, buf.size(), (int)(unsigned char)buf.front(), (int)(unsigned char)buf.back());
    printf("%u read %zu\n", k, split.BytesRead());
  }
  return 0;
}
