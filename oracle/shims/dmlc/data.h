// Stand-in for <dmlc/data.h>: the RowBlock view the reference's hot path consumes
// (dmlc-core not vendored). Field names/types follow the usage in
// /root/reference/learn/base/localizer.h:100-128,199-220 and spmv.h:72-119.
// Test infrastructure only.
#pragma once
#include <cstring>
#include "dmlc/base.h"
#include "dmlc/logging.h"
namespace dmlc {
template <typename IndexType>
struct Row {
  real_t label;
  real_t weight;
  size_t length;
  const IndexType* index;
  const real_t* value;
};
template <typename IndexType>
struct RowBlock {
  size_t size;
  const size_t* offset;
  const real_t* label;
  const real_t* weight;
  const IndexType* index;
  const real_t* value;
  RowBlock() : size(0), offset(NULL), label(NULL), weight(NULL), index(NULL), value(NULL) {}
  inline Row<IndexType> operator[](size_t i) const {
    Row<IndexType> r;
    r.label = label ? label[i] : 0;
    r.weight = weight ? weight[i] : 1;
    r.length = offset[i + 1] - offset[i];
    r.index = index + offset[i];
    r.value = value ? value + offset[i] : NULL;
    return r;
  }
};
}  // namespace dmlc
