/* CPU restatement of the reader in front of the hot path: the Criteo and libsvm text
 * parsers. TEST INFRASTRUCTURE ONLY -- the product parses on the GPU (wormhole_b200/csrc/parse.cu).
 *
 * Pinned against oracle/_ref/libwh_ref_io.so (the reference's own criteo_parser.h / minibatch_iter.h compiled
 * behind stand-ins) by tests/test_io_oracle.py. What stays unpinned is third-party: CityHash64 of
 * cityhash-1.1.1 (make/deps.mk:73-83, absent here; restated for <= 16 bytes, pinned only by the published value
 * for "abc") and dmlc-core's libsvm parser (cloned unpinned by the reference's Makefile:25-27; restated as
 * label + index[:value] tokens read with atof / strtoull).
 *
 * Malformed input: the reference's Criteo parser does not stop at line ends (its Find and its 8-byte
 * categorical read run on into the next line) and CHECK-fails on some of the results. Wherever it would leave
 * the line or abort, these functions return a non-zero code instead; on well-formed lines they follow it byte
 * for byte, including its habit of hashing 8 raw bytes of a categorical field whatever they contain. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define WH_O_API __attribute__((visibility("default")))

static uint64_t fetch64(const char* p) { uint64_t r; memcpy(&r, p, 8); return r; }
static uint32_t fetch32(const char* p) { uint32_t r; memcpy(&r, p, 4); return r; }
static uint64_t rot(uint64_t v, int s) { return s == 0 ? v : ((v >> s) | (v << (64 - s))); }
static uint64_t len16(uint64_t u, uint64_t v, uint64_t mul) {
  uint64_t a = (u ^ v) * mul, b;
  a ^= (a >> 47);
  b = (v ^ a) * mul;
  b ^= (b >> 47);
  return b * mul;
}
/* CityHash64 (cityhash-1.1.1 city.cc, HashLen0to16), len <= 16 */
WH_O_API uint64_t wh_o_cityhash64(const char* s, size_t len) {
  const uint64_t k0 = 0xc3a5c85c97cb3127ULL, k2 = 0x9ae16a3b2f90404fULL;
  if (len >= 8) {
    uint64_t mul = k2 + len * 2, a = fetch64(s) + k2, b = fetch64(s + len - 8);
    uint64_t c = rot(b, 37) * mul + a, d = (rot(a, 25) + b) * mul;
    return len16(c, d, mul);
  }
  if (len >= 4) {
    uint64_t mul = k2 + len * 2, a = fetch32(s);
    return len16(len + (a << 3), fetch32(s + len - 4), mul);
  }
  if (len > 0) {
    uint8_t a = (uint8_t)s[0], b = (uint8_t)s[len >> 1], c = (uint8_t)s[len - 1];
    uint32_t y = (uint32_t)a + ((uint32_t)b << 8), z = (uint32_t)len + ((uint32_t)c << 2);
    uint64_t m = y * k2 ^ z * k0;
    return (m ^ (m >> 47)) * k2;
  }
  return k2;
}

static int is_brk(char c) { return c == '\n' || c == '\r'; }
/* atof on a token that ends at `end` (the text need not be NUL-terminated) */
static double number(const char* p, const char* end) {
  char buf[64];
  size_t n = (size_t)(end - p);
  if (n > 63) n = 63;
  memcpy(buf, p, n);
  buf[n] = 0;
  return atof(buf);
}

/* learn/base/criteo_parser.h:38-89. Outputs sized by the caller: offset[lines+1], label[lines], index[39*lines].
 * Returns 0, or the code of the first malformed line (see the header comment) with *rows = its number. */
WH_O_API int wh_o_parse_criteo(const char* text, size_t len, int is_train, size_t* rows, size_t* nnz,
                               uint64_t* offset, float* label, uint64_t* index) {
  const char *p = text, *end = text + len;
  size_t r = 0, n = 0;
  int i;
  offset[0] = 0;
  while (p != end) {
    const char *lend, *pp;
    while (p != end && is_brk(*p)) ++p;                  /* :50 */
    if (p == end) break;
    lend = p;
    while (lend != end && !is_brk(*lend)) ++lend;
    *rows = r;
    if (is_train) {                                      /* :54-62 */
      pp = p;
      while (pp != lend && *pp != '\t') ++pp;
      if (pp == lend || pp == p) return 1;               /* no label: CHECK_NE / runs off the line */
      label[r] = (float)number(p, pp);
      p = pp + 1;
    } else {
      label[r] = 0;
    }
    for (i = 0; i < 13; ++i) {                           /* :65-72 */
      pp = p;
      while (pp != lend && *pp != '\t') ++pp;
      if (pp == lend) return 2;                          /* Find would leave the line */
      if (pp > p) {
        if (pp - p > 16) return 3;                       /* beyond the restated CityHash64 */
        index[n++] = (wh_o_cityhash64(p, (size_t)(pp - p)) >> 10) | ((uint64_t)i << 54);
      }
      p = pp + 1;
    }
    for (i = 0; i < 26; ++i) {                           /* :75-85 */
      if (p == end) break;
      if (p == lend) {                                   /* isspace(line break): swallowed as an empty field */
        if (i != 25) return 4;                           /* ... and the next line's bytes would follow */
        break;
      }
      if (*p == '\t') { ++p; continue; }
      if (*p == ' ' || *p == '\v' || *p == '\f') return 5;
      pp = p + 8;
      if (pp > lend) return 6;                           /* 8-byte read crosses the line end */
      if (pp != end && !(*pp == '\t' || is_brk(*pp) || *pp == ' ' || *pp == '\v' || *pp == '\f')) return 6; /* CHECK(isspace) */
      index[n++] = (wh_o_cityhash64(p, 8) >> 10) | ((uint64_t)(i + 13) << 54);
      p = pp == end ? end : pp + 1;
      if (pp == end || is_brk(*pp)) break;
    }
    if (i == 26 && p < lend) return 7;                   /* bytes left: the reference starts a bogus row there */
    p = lend;
    offset[++r] = n;                                     /* :86 */
  }
  *rows = r;
  *nnz = n;
  return 0;
}

/* libsvm text (dmlc-core LibSVMParser as restated in oracle/shims_io/data/libsvm_parser.h).
 * offset[lines+1], label[lines], index/value[tokens]. *nvalue = number of tokens that carried a value
 * (0: binary block, == *nnz: valued block, anything else is an error of the input). */
WH_O_API int wh_o_parse_libsvm(const char* text, size_t len, size_t* rows, size_t* nnz, size_t* nvalue,
                               uint64_t* offset, float* label, uint64_t* index, float* value) {
  const char *p = text, *end = text + len;
  size_t r = 0, n = 0, nv = 0;
  offset[0] = 0;
  while (p != end) {
    const char* lend = p;
    while (lend != end && !is_brk(*lend)) ++lend;
    while (p != lend && (*p == ' ' || *p == '\t')) ++p;
    if (p != lend) {
      const char* q = p;
      while (q != lend && *q != ' ' && *q != '\t') ++q;
      label[r] = (float)number(p, q);
      p = q;
      for (;;) {
        while (p != lend && (*p == ' ' || *p == '\t')) ++p;
        if (p == lend) break;
        q = p;
        while (q != lend && *q != ' ' && *q != '\t') ++q;
        {
          char buf[32];
          size_t m = (size_t)(q - p) < 31 ? (size_t)(q - p) : 31;
          const char* c = p;
          memcpy(buf, p, m);
          buf[m] = 0;
          index[n] = strtoull(buf, NULL, 10);
          while (c != q && *c != ':') ++c;
          if (c != q) { value[nv++] = (float)number(c + 1, q); }
        }
        ++n;
        p = q;
      }
      offset[++r] = n;
    }
    p = lend;
    while (p != end && is_brk(*p)) ++p;
  }
  *rows = r;
  *nnz = n;
  *nvalue = nv;
  return (nv == 0 || nv == n) ? 0 : 8;
}

/* learn/base/adfea_parser.h:38-82: a token stream; "idx:gid" tokens are features (index = idx >> 10 | gid << 54 for
 * 64-bit indices), every third other token opens a row whose label is (first character == '1'). Outputs sized by the
 * caller: offset[tokens+1], label[tokens], index[tokens]. Returns 0, or 10 where the reference's CHECK_NE(head, p)
 * fails (a token that does not start with a digit). */
static int is_sp(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\v' || c == '\f' || c == '\r'; }
WH_O_API int wh_o_parse_adfea(const char* text, size_t len, size_t* rows, size_t* nnz, uint64_t* offset, float* label,
                              uint64_t* index) {
  const char *p = text, *end = text + len;
  size_t r = 0, n = 0;
  int i = 0;
  offset[0] = 0;
  while (p != end && is_sp(*p)) ++p;                                   /* :48 */
  while (p != end) {
    const char* head = p;
    while (p != end && *p >= '0' && *p <= '9') ++p;
    if (head == p) return 10;                                          /* :53 CHECK_NE */
    if (p != end && *p == ':') {                                       /* :55-65 */
      char buf[32];
      size_t m = (size_t)(p - head) < 31 ? (size_t)(p - head) : 31;
      uint64_t idx, gid;
      const char* g;
      memcpy(buf, head, m); buf[m] = 0;
      idx = strtoull(buf, NULL, 10);
      g = ++p;
      while (p != end && *p >= '0' && *p <= '9') ++p;
      m = (size_t)(p - g) < 31 ? (size_t)(p - g) : 31;
      memcpy(buf, g, m); buf[m] = 0;
      gid = strtoull(buf, NULL, 10);
      index[n++] = (idx >> 10) | (gid << 54);
    } else {                                                           /* :67-77 */
      if (i == 2) {
        i = 0;
        if (r != 0) offset[r] = n;
        label[r++] = *head == '1' ? 1.f : 0.f;
      } else {
        ++i;
      }
    }
    while (p != end && is_sp(*p)) ++p;
  }
  if (r != 0) offset[r] = n;                                           /* :79-81 */
  *rows = r;
  *nnz = n;
  return 0;
}
