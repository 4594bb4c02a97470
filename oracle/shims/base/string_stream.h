// Stand-in for learn/base/string_stream.h inside oracle/_ref: an in-memory dmlc::Stream (append-only writes,
// sequential reads, str() returns everything written). It replaces the reference's header because the stand-in
// dmlc::Stream (shims/dmlc/io.h) also declares string overloads of Read / Write, which a derived class must
// re-export with using-declarations. Test infrastructure only.
#pragma once
#include <algorithm>
#include <string>
#include <vector>

#include "dmlc/io.h"

namespace dmlc {

class StringStream : public Stream {
 public:
  using Stream::Read;
  using Stream::Write;

  StringStream() {}
  explicit StringStream(const std::string& initial) : bytes_(initial.begin(), initial.end()) {}

  // reading past the end is a programming error in the callers (serialised Workload messages): abort like CHECK_LE
  size_t Read(void* dst, size_t n) override {
    CHECK(cursor_ + n <= bytes_.size()) << "StringStream: read of " << n << " bytes at " << cursor_ << " of " << bytes_.size();
    std::copy(bytes_.begin() + cursor_, bytes_.begin() + cursor_ + n, static_cast<char*>(dst));
    cursor_ += n;
    return n;
  }
  void Write(const void* src, size_t n) override {
    const char* p = static_cast<const char*>(src);
    bytes_.insert(bytes_.end(), p, p + n);
  }
  std::string str() const { return std::string(bytes_.begin(), bytes_.end()); }

 private:
  std::vector<char> bytes_;
  size_t cursor_ = 0;
};

}  // namespace dmlc
