// Shared plumbing of the C++ facade: one whContext per process and device, and the reference's
// error behaviour (glog CHECK / LOG(FATAL) abort the process; here a non-zero status from the
// C ABI prints the library's message and aborts).
#pragma once
#include <cstdio>
#include <cstdlib>

#include "wormhole_b200.h"

namespace wh {

inline void Check(int rc, const char* what) {
  if (rc != WH_OK) {
    fprintf(stderr, "[FATAL] %s: %s\n", what, wh_last_error());
    abort();
  }
}
#define WH_CHECK(call) ::wh::Check((call), #call)

// The process-wide context (device from $WH_DEVICE, default 0). The reference's primitives are
// free functions / stateless classes, so the facade needs an implicit home for the GPU state.
inline whContext* Ctx() {
  static whContext* ctx = nullptr;
  if (!ctx) {
    const char* d = getenv("WH_DEVICE");
    WH_CHECK(wh_create(d ? atoi(d) : 0, &ctx));
  }
  return ctx;
}

}  // namespace wh
