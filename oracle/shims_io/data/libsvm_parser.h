// Stand-in for dmlc-core's src/data/libsvm_parser.h (not vendored; the reference's build clones dmlc-core
// at its then-current head, Makefile:25-27, so no version is pinned). Restated from the format and from the
// 2015 parser's behaviour: one row per non-blank line, `label` then whitespace-separated `index[:value]`
// tokens; numbers by libc atof / strtoull, cast to real_t / IndexType. A row block has values iff every
// feature token carries one. Parity: unpinned. Test infrastructure only.
#pragma once
#include <ctype.h>
#include <stdlib.h>
#include "data/parser.h"
namespace dmlc {
namespace data {
template <typename IndexType>
class LibSVMParser : public ParserImpl<IndexType> {
 public:
  explicit LibSVMParser(InputSplit* source, int /*nthread*/) : bytes_read_(0), source_(source) {}
  virtual ~LibSVMParser() { delete source_; }
  virtual void BeforeFirst(void) { source_->BeforeFirst(); }
  virtual size_t BytesRead(void) const { return bytes_read_; }
  virtual bool ParseNext(std::vector<RowBlockContainer<IndexType> >* data) {
    InputSplit::Blob chunk;
    if (!source_->NextChunk(&chunk)) return false;
    bytes_read_ += chunk.size;
    data->resize(1);
    RowBlockContainer<IndexType>& blk = (*data)[0];
    blk.Clear();
    const char* p = reinterpret_cast<const char*>(chunk.dptr);
    const char* end = p + chunk.size;
    while (p != end) {
      const char* lend = p;
      while (lend != end && *lend != '\n' && *lend != '\r') ++lend;
      while (p != lend && (*p == ' ' || *p == '\t')) ++p;
      if (p != lend) {
        blk.label.push_back(static_cast<real_t>(atof(p)));
        while (p != lend && *p != ' ' && *p != '\t') ++p;
        while (true) {
          while (p != lend && (*p == ' ' || *p == '\t')) ++p;
          if (p == lend) break;
          blk.index.push_back(static_cast<IndexType>(strtoull(p, NULL, 10)));
          while (p != lend && *p != ' ' && *p != '\t' && *p != ':') ++p;
          if (p != lend && *p == ':') {
            blk.value.push_back(static_cast<real_t>(atof(p + 1)));
            while (p != lend && *p != ' ' && *p != '\t') ++p;
          }
        }
        blk.offset.push_back(blk.index.size());
      }
      p = lend;
      while (p != end && (*p == '\n' || *p == '\r')) ++p;
    }
    CHECK(blk.value.empty() || blk.value.size() == blk.index.size()) << "some features lack a value";
    return true;
  }
 private:
  size_t bytes_read_;
  InputSplit* source_;
};
}  // namespace data
}  // namespace dmlc
