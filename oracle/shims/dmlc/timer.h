// Stand-in for <dmlc/timer.h> (dmlc-core not vendored). Test infrastructure only.
#pragma once
#include <chrono>
namespace dmlc {
inline double GetTime() {
  return std::chrono::duration<double>(
      std::chrono::steady_clock::now().time_since_epoch()).count();
}
}  // namespace dmlc
