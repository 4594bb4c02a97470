// Minimal dmlc::RowBlock for builds without dmlc-core (which the reference fetches at build time).
// With dmlc-core on the include path, use its <dmlc/data.h> instead: field names and types match
// what learn/base/localizer.h:100-128 and learn/base/spmv.h:72-119 touch.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>
namespace dmlc {
typedef float real_t;
template <typename IndexType>
struct RowBlock {
  size_t size = 0;
  const size_t* offset = nullptr;
  const real_t* label = nullptr;
  const real_t* weight = nullptr;
  const IndexType* index = nullptr;
  const real_t* value = nullptr;
};
template <typename T> inline T* BeginPtr(std::vector<T>& v) { return v.empty() ? nullptr : v.data(); }
template <typename T> inline const T* BeginPtr(const std::vector<T>& v) { return v.empty() ? nullptr : v.data(); }
}  // namespace dmlc
