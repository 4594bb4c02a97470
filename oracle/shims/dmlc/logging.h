// Stand-in for dmlc-core's <dmlc/logging.h> (dmlc-core is not vendored in the
// reference tree: /root/reference/Makefile:25 clones it at build time).
// TEST INFRASTRUCTURE ONLY: lets the reference's own headers under
// /root/reference/learn compile unchanged into oracle/_ref/. Not product code.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <sstream>
#include <string>

namespace dmlc {
class LogMessageFatal {
 public:
  LogMessageFatal(const char* file, int line) { s_ << file << ":" << line << ": "; }
  std::ostringstream& stream() { return s_; }
  ~LogMessageFatal() {
    fprintf(stderr, "[FATAL] %s\n", s_.str().c_str());
    abort();
  }
 private:
  std::ostringstream s_;
};
class LogMessageQuiet {
 public:
  LogMessageQuiet(const char* file, int line, bool show) : show_(show) {
    if (show_) s_ << file << ":" << line << ": ";
  }
  std::ostringstream& stream() { return s_; }
  ~LogMessageQuiet() { if (show_) fprintf(stderr, "%s\n", s_.str().c_str()); }
 private:
  bool show_;
  std::ostringstream s_;
};
inline bool ShowInfoLog() {
  static int v = -1;
  if (v < 0) v = getenv("WH_ORACLE_VERBOSE") ? 1 : 0;
  return v == 1;
}
template <typename T> inline T* CheckNotNullImpl(const char* f, int l, const char* what, T* p) {
  if (p == NULL) { LogMessageFatal(f, l).stream() << "Check notnull: " << what; }
  return p;
}
}  // namespace dmlc

#define LOG_INFO dmlc::LogMessageQuiet(__FILE__, __LINE__, dmlc::ShowInfoLog())
#define LOG_WARNING dmlc::LogMessageQuiet(__FILE__, __LINE__, true)
#define LOG_ERROR dmlc::LogMessageQuiet(__FILE__, __LINE__, true)
#define LOG_FATAL dmlc::LogMessageFatal(__FILE__, __LINE__)
#define LOG(severity) LOG_##severity.stream()
#define LL LOG_ERROR.stream()

#define CHECK(x) \
  if (!(x)) dmlc::LogMessageFatal(__FILE__, __LINE__).stream() << "Check failed: " #x << ' '
#define CHECK_BINARY_(a, b, op) \
  if (!((a) op (b))) dmlc::LogMessageFatal(__FILE__, __LINE__).stream() \
      << "Check failed: " #a " " #op " " #b << ' '
#define CHECK_LT(a, b) CHECK_BINARY_(a, b, <)
#define CHECK_GT(a, b) CHECK_BINARY_(a, b, >)
#define CHECK_LE(a, b) CHECK_BINARY_(a, b, <=)
#define CHECK_GE(a, b) CHECK_BINARY_(a, b, >=)
#define CHECK_EQ(a, b) CHECK_BINARY_(a, b, ==)
#define CHECK_NE(a, b) CHECK_BINARY_(a, b, !=)
#define CHECK_NOTNULL(x) dmlc::CheckNotNullImpl(__FILE__, __LINE__, #x, (x))
