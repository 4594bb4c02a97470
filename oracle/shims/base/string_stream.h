// Shadows /root/reference/learn/base/string_stream.h so that oracle/_ref does not
// depend on which dmlc::Stream::Write signature is in scope. Same behaviour:
// an in-memory Stream. Test infrastructure only.
#pragma once
#include <cstring>
#include <string>
#include "dmlc/io.h"
namespace dmlc {
class StringStream : public Stream {
 public:
  StringStream() : pos_(0) {}
  explicit StringStream(const std::string& s) : buf_(s), pos_(0) {}
  using Stream::Read;
  using Stream::Write;
  virtual size_t Read(void* ptr, size_t size) {
    CHECK_LE(size + pos_, buf_.size());
    memcpy(ptr, buf_.data() + pos_, size); pos_ += size; return size;
  }
  virtual void Write(const void* ptr, size_t size) { buf_.append((const char*)ptr, size); }
  std::string str() { return buf_; }
 private:
  std::string buf_;
  size_t pos_;
};
}  // namespace dmlc
