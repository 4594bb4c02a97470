// Shadows /root/reference/learn/base/minibatch_iter.h (needs dmlc-core parsers,
// cityhash, lz4: none vendored). The oracle harness feeds RowBlock minibatches
// directly into ProcessMinibatch, so the reader is never run. Test infrastructure only.
#pragma once
#include "data/row_block.h"
namespace dmlc {
namespace data {
template <typename IndexType>
class MinibatchIter {
 public:
  MinibatchIter(const char*, unsigned, unsigned, const char*, unsigned, unsigned = 0, float = 1.0) {}
  void BeforeFirst() {}
  bool Next() { return false; }
  const RowBlock<IndexType>& Value() const { return blk_; }
 private:
  RowBlock<IndexType> blk_;
};
}  // namespace data
}  // namespace dmlc
