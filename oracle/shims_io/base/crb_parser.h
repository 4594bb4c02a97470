// Shadows learn/base/crb_parser.h: the compressed row block format needs lz4 and dmlc recordio, neither vendored
// nor in this image. Opening a "crb" source aborts. Test infrastructure only.
#pragma once
#include "data/parser.h"
namespace dmlc {
namespace data {
template <typename IndexType>
class CRBParser : public ParserImpl<IndexType> {
 public:
  explicit CRBParser(InputSplit*) { LOG(FATAL) << "crb: lz4/recordio not available"; }
  virtual void BeforeFirst(void) {}
  virtual size_t BytesRead(void) const { return 0; }
  virtual bool ParseNext(std::vector<RowBlockContainer<IndexType> >*) { return false; }
};
}  // namespace data
}  // namespace dmlc
