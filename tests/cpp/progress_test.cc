// Prints the progress-table header and rows of the facade's Progress structs for vectors read from stdin
// ("L v0..v5" or "D v0..v7" per line); tests/test_solver_host.py compares them with the reference's own output.
#include <cstdio>
#include <cstring>

#include "difacto/progress.h"
#include "linear/progress.h"

int main() {
  dmlc::linear::Progress lin;
  dmlc::difacto::Progress dif;
  printf("%s\n%s\n", dmlc::linear::Progress::HeadStr().c_str(), dmlc::difacto::Progress::HeadStr().c_str());
  char kind;
  while (scanf(" %c", &kind) == 1) {
    if (kind == 'L') {
      for (int i = 0; i < 6; ++i) if (scanf("%lf", &lin.data[i]) != 1) return 1;
      printf("%s\n", lin.PrintStr().c_str());
    } else {
      for (int i = 0; i < 8; ++i) if (scanf("%lf", &dif.data[i]) != 1) return 1;
      printf("%s\n", dif.PrintStr().c_str());
    }
  }
  return 0;
}
