// Stand-in for dmlc-core's src/data/row_block.h (RowBlockContainer), which the
// reference includes as "data/row_block.h" (/root/reference/learn/base/localizer.h:6)
// but does not vendor. Test infrastructure only.
#pragma once
#include <algorithm>
#include <vector>
#include "dmlc/data.h"
namespace dmlc {
namespace data {
template <typename IndexType>
struct RowBlockContainer {
  std::vector<size_t> offset;
  std::vector<real_t> label;
  std::vector<real_t> weight;
  std::vector<IndexType> index;
  std::vector<real_t> value;
  IndexType max_index;
  RowBlockContainer() { Clear(); }
  inline void Clear() {
    offset.clear(); offset.push_back(0);
    label.clear(); weight.clear(); index.clear(); value.clear();
    max_index = 0;
  }
  inline size_t Size() const { return offset.size() - 1; }
  inline RowBlock<IndexType> GetBlock() const {
    RowBlock<IndexType> d;
    d.size = offset.size() - 1;
    d.offset = BeginPtr(offset);
    d.label = BeginPtr(label);
    d.weight = BeginPtr(weight);
    d.index = BeginPtr(index);
    d.value = BeginPtr(value);
    return d;
  }
  template <typename I>
  inline void Push(Row<I> row) {
    label.push_back(row.label);
    for (size_t i = 0; i < row.length; ++i) {
      index.push_back(static_cast<IndexType>(row.index[i]));
      max_index = std::max(max_index, index.back());
    }
    if (row.value) for (size_t i = 0; i < row.length; ++i) value.push_back(row.value[i]);
    offset.push_back(index.size());
  }
  template <typename I>
  inline void Push(RowBlock<I> b) {
    size_t n = b.offset[b.size] - b.offset[0];
    for (size_t i = 0; i < b.size; ++i) label.push_back(b.label[i]);
    for (size_t i = 0; i < n; ++i) {
      index.push_back(static_cast<IndexType>(b.index[i]));
      max_index = std::max(max_index, index.back());
    }
    if (b.value) for (size_t i = 0; i < n; ++i) value.push_back(b.value[i]);
    size_t shift = offset.back();
    for (size_t i = 0; i < b.size; ++i) offset.push_back(shift + b.offset[i + 1] - b.offset[0]);
  }
};
}  // namespace data
}  // namespace dmlc
