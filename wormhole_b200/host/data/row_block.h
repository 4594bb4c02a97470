// Minimal dmlc::data::RowBlockContainer (dmlc-core src/data/row_block.h), see dmlc/data.h here.
#pragma once
#include "dmlc/data.h"
namespace dmlc {
namespace data {
template <typename IndexType>
struct RowBlockContainer {
  std::vector<size_t> offset;
  std::vector<real_t> label, weight;
  std::vector<IndexType> index;
  std::vector<real_t> value;
  IndexType max_index = 0;
  RowBlockContainer() { Clear(); }
  void Clear() { offset.assign(1, 0); label.clear(); weight.clear(); index.clear(); value.clear(); max_index = 0; }
  size_t Size() const { return offset.size() - 1; }
  RowBlock<IndexType> GetBlock() const {
    RowBlock<IndexType> d;
    d.size = offset.size() - 1; d.offset = BeginPtr(offset); d.label = BeginPtr(label);
    d.weight = BeginPtr(weight); d.index = BeginPtr(index); d.value = BeginPtr(value);
    return d;
  }
};
}  // namespace data
}  // namespace dmlc
