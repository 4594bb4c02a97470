// Shadows /root/reference/learn/base/match_file.h (needs dmlc-core io/filesys.h,
// not vendored). File matching is control plane, outside the hot path.
// Test infrastructure only.
#pragma once
#include <string>
#include <vector>
namespace dmlc {
inline void MatchFile(const std::string& pattern, std::vector<std::string>* matched) {
  matched->push_back(pattern);
}
}  // namespace dmlc
