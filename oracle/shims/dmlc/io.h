// Stand-in for <dmlc/io.h> (dmlc-core not vendored). Only what the reference's
// solver/worker headers touch: Stream, Serializable, dmlc::ostream.
// Test infrastructure only.
#pragma once
#include <cstring>
#include <ostream>
#include <streambuf>
#include <string>
#include <vector>
#include "dmlc/base.h"
#include "dmlc/logging.h"
namespace dmlc {
class Stream {
 public:
  virtual size_t Read(void* ptr, size_t size) = 0;
  virtual void Write(const void* ptr, size_t size) = 0;
  virtual ~Stream() {}
  static Stream* Create(const char* uri, const char* const flag, bool allow_null = false);
  void Write(const std::string& s) {
    uint64_t n = s.size(); Write(&n, sizeof(n)); if (n) Write(s.data(), n);
  }
  bool Read(std::string* s) {
    uint64_t n; if (Read(&n, sizeof(n)) != sizeof(n)) return false;
    s->resize(n); if (n) return Read(&(*s)[0], n) == n; return true;
  }
};
class Serializable {
 public:
  virtual ~Serializable() {}
  virtual void Load(Stream* fi) = 0;
  virtual void Save(Stream* fo) const = 0;
};
class ostream : public std::basic_ostream<char> {
 public:
  explicit ostream(Stream* stream, size_t = (1 << 10)) : std::basic_ostream<char>(NULL), buf_(stream) {
    this->rdbuf(&buf_);
  }
  virtual ~ostream() { buf_.pubsync(); }
 private:
  class OutBuf : public std::streambuf {
   public:
    explicit OutBuf(Stream* s) : s_(s) {}
   protected:
    int_type overflow(int_type c) override {
      if (c != traits_type::eof()) { char ch = static_cast<char>(c); s_->Write(&ch, 1); }
      return c;
    }
    std::streamsize xsputn(const char* p, std::streamsize n) override { s_->Write(p, n); return n; }
   private:
    Stream* s_;
  };
  OutBuf buf_;
};
}  // namespace dmlc
