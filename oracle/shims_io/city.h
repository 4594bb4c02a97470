// Stand-in for <city.h> of cityhash-1.1.1 (make/deps.mk:73-83; third-party, not vendored, not in this image).
// CityHash64 restated from the published v1.1 algorithm for inputs of at most 16 bytes -- all the reference's
// Criteo parser ever hashes on well-formed data (integer fields and 8-character categorical fields,
// learn/base/criteo_parser.h:66-84). Longer inputs abort. Pinned only by the published value
// CityHash64("abc") = 2640714258260161385 (tests/test_io_oracle.py). Test infrastructure only.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
typedef uint64_t uint64;
static inline uint64_t wh_city_fetch64(const char* p) { uint64_t r; memcpy(&r, p, 8); return r; }
static inline uint32_t wh_city_fetch32(const char* p) { uint32_t r; memcpy(&r, p, 4); return r; }
static inline uint64_t wh_city_rot(uint64_t v, int s) { return s == 0 ? v : ((v >> s) | (v << (64 - s))); }
static inline uint64_t wh_city_len16(uint64_t u, uint64_t v, uint64_t mul) {
  uint64_t a = (u ^ v) * mul; a ^= (a >> 47);
  uint64_t b = (v ^ a) * mul; b ^= (b >> 47);
  return b * mul;
}
static inline uint64_t CityHash64(const char* s, size_t len) {
  const uint64_t k0 = 0xc3a5c85c97cb3127ULL, k2 = 0x9ae16a3b2f90404fULL;
  if (len > 16) abort();
  if (len >= 8) {
    uint64_t mul = k2 + len * 2, a = wh_city_fetch64(s) + k2, b = wh_city_fetch64(s + len - 8);
    uint64_t c = wh_city_rot(b, 37) * mul + a, d = (wh_city_rot(a, 25) + b) * mul;
    return wh_city_len16(c, d, mul);
  }
  if (len >= 4) {
    uint64_t mul = k2 + len * 2, a = wh_city_fetch32(s);
    return wh_city_len16(len + (a << 3), wh_city_fetch32(s + len - 4), mul);
  }
  if (len > 0) {
    uint8_t a = (uint8_t)s[0], b = (uint8_t)s[len >> 1], c = (uint8_t)s[len - 1];
    uint32_t y = (uint32_t)a + ((uint32_t)b << 8), z = (uint32_t)len + ((uint32_t)c << 2);
    uint64_t m = y * k2 ^ z * k0;
    return (m ^ (m >> 47)) * k2;
  }
  return k2;
}
