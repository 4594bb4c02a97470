// Stand-in for <dmlc/base.h> (dmlc-core not vendored). Test infrastructure only.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>
namespace dmlc {
typedef float real_t;
template <typename T> inline T* BeginPtr(std::vector<T>& v) { return v.empty() ? NULL : &v[0]; }
template <typename T> inline const T* BeginPtr(const std::vector<T>& v) { return v.empty() ? NULL : &v[0]; }
}  // namespace dmlc
