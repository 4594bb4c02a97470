// Stand-in for dmlc-core's src/data/parser.h + <dmlc/io.h> InputSplit (not vendored): ParserImpl drives
// ParseNext and hands out one RowBlock per parsed chunk; ThreadedParser is the same thing without the
// prefetch thread; InputSplit("text") serves byte range `part` of `nparts` of one file, extended to whole
// lines (a part starts after the first line break at or after its byte offset, dmlc-core's rule), in
// chunks of WH_ORACLE_CHUNK bytes (default 8 MiB) cut at line ends. Test infrastructure only.
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include <vector>
#include "dmlc/data.h"
#include "dmlc/logging.h"
#include "data/row_block.h"
namespace dmlc {
class InputSplit {
 public:
  struct Blob { void* dptr; size_t size; };
  static InputSplit* Create(const char* uri, unsigned part, unsigned nparts, const char* type) {
    CHECK(std::string(type) == "text") << "only text splits";
    return new InputSplit(uri, part, nparts);
  }
  void BeforeFirst() { pos_ = 0; }
  bool NextChunk(Blob* out) {
    if (pos_ >= size_) return false;
    size_t end = pos_ + chunk_;
    if (end >= size_) end = size_;
    else {
      while (end < size_ && !brk(buf_[end])) ++end;   // finish the line
      while (end < size_ && brk(buf_[end])) ++end;    // and its break
    }
    out->dptr = &buf_[pos_]; out->size = end - pos_; pos_ = end;
    return true;
  }
 private:
  static bool brk(char c) { return c == '\n' || c == '\r'; }
  InputSplit(const char* uri, unsigned part, unsigned nparts) : pos_(0) {
    const char* e = getenv("WH_ORACLE_CHUNK");
    chunk_ = e ? (size_t)atoll(e) : (size_t)8 << 20;
    FILE* f = fopen(uri, "rb"); CHECK(f != NULL) << "cannot open " << uri;
    fseek(f, 0, SEEK_END); size_t total = ftell(f); fseek(f, 0, SEEK_SET);
    std::vector<char> all(total + 1);
    if (total) CHECK(fread(all.data(), 1, total, f) == total);
    fclose(f); all[total] = 0;
    size_t b = total * part / nparts, e2 = total * (part + 1) / nparts;
    if (part) { while (b < total && !brk(all[b])) ++b; while (b < total && brk(all[b])) ++b; }
    if (part + 1 < nparts) { while (e2 < total && !brk(all[e2])) ++e2; while (e2 < total && brk(all[e2])) ++e2; }
    if (e2 < b) e2 = b;
    buf_.assign(all.begin() + b, all.begin() + e2);
    size_ = buf_.size();
    buf_.push_back(0);           // the parsers peek one byte past a chunk's end
  }
  std::vector<char> buf_;        // size_ bytes of text + a terminating 0
  size_t size_, pos_, chunk_;
};
namespace data {
template <typename IndexType>
class ParserImpl {
 public:
  ParserImpl() : ptr_(0), end_(0) {}
  virtual ~ParserImpl() {}
  virtual bool Next(void) {
    while (true) {
      while (ptr_ < end_) { ptr_ += 1; if (data_[ptr_ - 1].Size() != 0) { block_ = data_[ptr_ - 1].GetBlock(); return true; } }
      if (!ParseNext(&data_)) break;
      ptr_ = 0; end_ = data_.size();
    }
    return false;
  }
  virtual const RowBlock<IndexType>& Value(void) const { return block_; }
  virtual size_t BytesRead(void) const = 0;
  virtual void BeforeFirst(void) = 0;
  virtual bool ParseNext(std::vector<RowBlockContainer<IndexType> >* data) = 0;
 protected:
  size_t ptr_, end_;
  std::vector<RowBlockContainer<IndexType> > data_;
  RowBlock<IndexType> block_;
};
template <typename IndexType>
class ThreadedParser : public ParserImpl<IndexType> {
 public:
  explicit ThreadedParser(ParserImpl<IndexType>* base) : base_(base) {}
  virtual ~ThreadedParser() { delete base_; }
  virtual void BeforeFirst(void) { base_->BeforeFirst(); this->ptr_ = this->end_ = 0; }
  virtual size_t BytesRead(void) const { return base_->BytesRead(); }
  virtual bool ParseNext(std::vector<RowBlockContainer<IndexType> >* data) { return base_->ParseNext(data); }
 private:
  ParserImpl<IndexType>* base_;
};
}  // namespace data
}  // namespace dmlc
