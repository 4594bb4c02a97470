// The reader in front of the hot path, on the GPU: raw text goes over PCIe (a Criteo line is
// ~230 bytes, its parsed row 312) and is turned into a device-resident dmlc::RowBlock<uint64_t> here.
//   wh_parse_text     learn/base/criteo_parser.h:38-89 (criteo, criteo_test); dmlc-core LibSVMParser (libsvm)
//   wh_rows_*         data::RowBlockContainer as MinibatchIter uses it (learn/base/minibatch_iter.h:74-147)
// One thread walks one line. Line starts come from a scan over "first byte after a line break" flags; row
// lengths from a second scan; nothing is parsed on the host.
#include <vector>

#include "common.cuh"
#include "scan.cuh"

struct whRows {
  whContext* ctx = nullptr;
  size_t rows = 0, nnz = 0;
  bool has_value = false;
  wh::DevBuf offset;   // size_t[rows+1]
  wh::DevBuf label;    // float[rows]
  wh::DevBuf index;    // uint64[nnz]
  wh::DevBuf value;    // float[nnz]
  std::vector<size_t> h_offset;   // host mirror of offset (lazy; valid when h_valid)
  bool h_valid = false;
  ~whRows() { offset.release(); label.release(); index.release(); value.release(); }
};

namespace wh {
namespace {

constexpr int kCriteoInts = 13, kCriteoCats = 26, kCriteoMax = kCriteoInts + kCriteoCats;

__device__ __forceinline__ bool is_brk(char c) { return c == '\n' || c == '\r'; }
__device__ __forceinline__ bool is_blank(char c) { return c == ' ' || c == '\t'; }
__device__ __forceinline__ bool is_digit(char c) { return c >= '0' && c <= '9'; }

// ---- CityHash64 for <= 16 bytes (cityhash-1.1.1 city.cc HashLen0to16; see oracle/wh_oracle_io.c) ----------
__device__ __forceinline__ uint64_t rot64(uint64_t v, int s) { return (v >> s) | (v << (64 - s)); }
__device__ __forceinline__ uint64_t city_len16(uint64_t u, uint64_t v, uint64_t mul) {
  uint64_t a = (u ^ v) * mul;
  a ^= (a >> 47);
  uint64_t b = (v ^ a) * mul;
  b ^= (b >> 47);
  return b * mul;
}
__device__ __forceinline__ uint64_t load_le(const char* p, int n) {   // n <= 8 bytes, little endian
  uint64_t r = 0;
  for (int i = 0; i < n; ++i) r |= (uint64_t)(uint8_t)p[i] << (8 * i);
  return r;
}
__device__ uint64_t cityhash64_short(const char* s, int len) {
  const uint64_t k0 = 0xc3a5c85c97cb3127ULL, k2 = 0x9ae16a3b2f90404fULL;
  if (len >= 8) {
    const uint64_t mul = k2 + (uint64_t)len * 2, a = load_le(s, 8) + k2, b = load_le(s + len - 8, 8);
    const uint64_t c = rot64(b, 37) * mul + a, d = (rot64(a, 25) + b) * mul;
    return city_len16(c, d, mul);
  }
  if (len >= 4) {
    const uint64_t mul = k2 + (uint64_t)len * 2, a = load_le(s, 4);
    return city_len16((uint64_t)len + (a << 3), load_le(s + len - 4, 4), mul);
  }
  if (len > 0) {
    const uint8_t a = (uint8_t)s[0], b = (uint8_t)s[len >> 1], c = (uint8_t)s[len - 1];
    const uint32_t y = (uint32_t)a + ((uint32_t)b << 8), z = (uint32_t)len + ((uint32_t)c << 2);
    const uint64_t m = y * k2 ^ z * k0;
    return (m ^ (m >> 47)) * k2;
  }
  return k2;
}

// ---- numbers ------------------------------------------------------------------------------------------------
// atof on [p, end): sign, digits, fraction, exponent; anything after the longest valid prefix is ignored and no
// digits at all give 0, as strtod does. Fast path: the value is formed as M * 10^e with ONE correctly rounded double
// operation, which equals strtod whenever M < 2^53 and |e| <= 22. Everything else -- 16 and more significant digits
// (repr(float), "%.17g"), exponents beyond +-22 -- takes the exact path below: big-integer arithmetic on up to 36
// significant digits (further ones only feed the sticky bit), rounded to nearest-even once, like strtod. Every caller
// stores the result as float, so magnitudes beyond the float range return +-inf and below it 0 without the long
// arithmetic. Only inf / nan / hex literals set *bad: there the reference's behaviour is genuinely different.
__constant__ double kP10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11,
                                1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
constexpr int kBigLimbs = 20;   // 640 bits: 36 digits (120 bits) x 10^40, or shifted past a 10^86 divisor
struct Big { uint32_t w[kBigLimbs]; };
__device__ void big_zero(Big& b) { for (int i = 0; i < kBigLimbs; ++i) b.w[i] = 0; }
__device__ void big_mul_add(Big& b, uint32_t m, uint32_t add) {
  uint64_t c = add;
  for (int i = 0; i < kBigLimbs; ++i) { c += (uint64_t)b.w[i] * m; b.w[i] = (uint32_t)c; c >>= 32; }
}
__device__ int big_bits(const Big& b) {
  for (int i = kBigLimbs - 1; i >= 0; --i) if (b.w[i]) return 32 * i + 32 - __clz(b.w[i]);
  return 0;
}
__device__ int big_bit(const Big& b, int i) { return (b.w[i >> 5] >> (i & 31)) & 1; }
__device__ void big_shl1_or(Big& b, uint32_t bit) {
  for (int i = kBigLimbs - 1; i > 0; --i) b.w[i] = (b.w[i] << 1) | (b.w[i - 1] >> 31);
  b.w[0] = (b.w[0] << 1) | bit;
}
__device__ bool big_ge(const Big& a, const Big& b) {
  for (int i = kBigLimbs - 1; i >= 0; --i) if (a.w[i] != b.w[i]) return a.w[i] > b.w[i];
  return true;
}
__device__ void big_sub(Big& a, const Big& b) {
  uint64_t br = 0;
  for (int i = 0; i < kBigLimbs; ++i) { const uint64_t d = (uint64_t)a.w[i] - b.w[i] - br; a.w[i] = (uint32_t)d; br = (d >> 32) & 1; }
}
// bits [lo, lo + 64) of b
__device__ uint64_t big_extract64(const Big& b, int lo) {
  uint64_t v = 0;
  for (int i = 0; i < 64; ++i) { const int k = lo + i; if (k >= 0 && k < 32 * kBigLimbs) v |= (uint64_t)big_bit(b, k) << i; }
  return v;
}
__device__ bool big_any_below(const Big& b, int hi) {
  for (int k = 0; k < hi && k < 32 * kBigLimbs; ++k) if (big_bit(b, k)) return true;
  return false;
}
// the double nearest to digits x 10^e10 (ties to even); digits = up to 36 decimal digits in `m`, sticky = a non-zero tail was dropped
__device__ double exact_decimal(const Big& m, int e10, bool sticky) {
  Big q;
  int exp2 = 0;
  if (e10 >= 0) {
    q = m;
    for (int i = 0; i < e10; ++i) big_mul_add(q, 10u, 0u);
  } else {
    Big den, rem;
    big_zero(den); den.w[0] = 1;
    for (int i = 0; i < -e10; ++i) big_mul_add(den, 10u, 0u);
    const int s = max(0, big_bits(den) + 66 - big_bits(m));
    // num = m << s, divided by den bit by bit (restoring division); the quotient has 66 or 67 bits
    big_zero(q); big_zero(rem);
    const int nb = big_bits(m) + s;
    for (int i = nb - 1; i >= 0; --i) {
      const int src = i - s;
      big_shl1_or(rem, src >= 0 ? (uint32_t)big_bit(m, src) : 0u);
      const bool ge = big_ge(rem, den);
      if (ge) big_sub(rem, den);
      big_shl1_or(q, ge ? 1u : 0u);
    }
    if (big_bits(rem)) sticky = true;
    exp2 = -s;
  }
  const int L = big_bits(q);
  if (L <= 53) return ldexp((double)big_extract64(q, 0), exp2);   // exact (sticky can only be set with a long quotient)
  const int drop = L - 53;
  uint64_t mant = big_extract64(q, drop) & ((1ULL << 53) - 1);
  const bool guard = big_bit(q, drop - 1) != 0;
  const bool rest = sticky || big_any_below(q, drop - 1);
  if (guard && (rest || (mant & 1))) ++mant;     // may reach 2^53: still exactly representable
  return ldexp((double)mant, exp2 + drop);
}

__device__ double parse_number(const char* p, const char* end, int* bad) {
  while (p != end && (is_blank(*p) || *p == '\v' || *p == '\f')) ++p;
  bool neg = false;
  if (p != end && (*p == '-' || *p == '+')) { neg = *p == '-'; ++p; }
  if (p != end && (*p == 'i' || *p == 'I' || *p == 'n' || *p == 'N')) { *bad = 1; return 0; }
  if (p + 1 < end && *p == '0' && (p[1] == 'x' || p[1] == 'X')) { *bad = 1; return 0; }
  const char* digits = p;
  uint64_t m = 0;
  int e = 0, ndig = 0;
  bool any = false, big = false;
  for (; p != end && is_digit(*p); ++p) {
    any = true;
    if (m == 0 && *p == '0') continue;
    if (++ndig > 18) { big = true; ++e; } else m = m * 10 + (uint64_t)(*p - '0');
  }
  if (p != end && *p == '.') {
    ++p;
    for (; p != end && is_digit(*p); ++p) {
      any = true;
      if (m == 0 && *p == '0') { --e; continue; }
      if (++ndig > 18) big = true; else { m = m * 10 + (uint64_t)(*p - '0'); --e; }
    }
  }
  if (!any) return 0;
  const char* digits_end = p;
  int ex = 0;
  if (p != end && (*p == 'e' || *p == 'E')) {
    const char* q = p + 1;
    bool eneg = false;
    if (q != end && (*q == '-' || *q == '+')) { eneg = *q == '-'; ++q; }
    if (q != end && is_digit(*q)) {
      int x = 0;
      for (; q != end && is_digit(*q); ++q) x = x < 10000 ? x * 10 + (*q - '0') : x;
      ex = eneg ? -x : x;
    }
  }
  double r;
  if (m == 0) r = 0;
  else if (!big && m < (1ULL << 53) && e + ex <= 22 && e + ex >= -22) {
    const int ee = e + ex;
    r = ee >= 0 ? (double)m * kP10[ee] : (double)m / kP10[-ee];
  } else {
    // the exact path: walk the digits again, keeping 36 of them
    Big M;
    big_zero(M);
    int kept = 0, e10 = ex;
    bool sticky = false, frac = false, lead = true;
    for (const char* d = digits; d != digits_end; ++d) {
      if (*d == '.') { frac = true; continue; }
      const uint32_t v = (uint32_t)(*d - '0');
      if (lead && v == 0) { if (frac) --e10; continue; }
      lead = false;
      if (kept < 36) { big_mul_add(M, 10u, v); ++kept; if (frac) --e10; }
      else { if (v) sticky = true; if (!frac) ++e10; }
    }
    const int mag = kept + e10;                 // value in [10^(mag-1), 10^mag)
    if (mag > 40) r = HUGE_VAL;                 // beyond the float range either way: the stored float is inf
    else if (mag < -50) r = 0;                  // ... or 0
    else r = exact_decimal(M, e10, sticky);
  }
  return neg ? -r : r;
}

// strtoull(p, NULL, 10) on [p, end)
__device__ uint64_t parse_u64(const char* p, const char* end) {
  while (p != end && (is_blank(*p) || *p == '\v' || *p == '\f')) ++p;
  bool neg = false;
  if (p != end && (*p == '-' || *p == '+')) { neg = *p == '-'; ++p; }
  uint64_t v = 0;
  bool over = false;
  for (; p != end && is_digit(*p); ++p) {
    const uint64_t d = (uint64_t)(*p - '0');
    if (v > (UINT64_MAX - d) / 10) over = true; else v = v * 10 + d;
  }
  if (over) return UINT64_MAX;
  return neg ? (uint64_t)0 - v : v;
}

// ---- lines --------------------------------------------------------------------------------------------------
// 1 where a line starts: a byte that is not a line break and follows one (or starts the text)
struct LineStartFlag {
  const char* t;
  __device__ unsigned operator()(size_t i) const { return (!is_brk(t[i]) && (i == 0 || is_brk(t[i - 1]))) ? 1u : 0u; }
};
struct EmitLineStart {
  uint32_t* starts;
  __device__ void operator()(size_t i, unsigned v, unsigned prefix) const { if (v) starts[prefix] = (uint32_t)i; }
};
__device__ __forceinline__ const char* line_end(const char* p, const char* end) {
  while (p != end && !is_brk(*p)) ++p;
  return p;
}
__device__ __forceinline__ void report(unsigned long long* err, int code, size_t line) {
  // keep the failure of the earliest line: (line+1) in the high bits decides under atomicMin
  atomicMin(err, ((unsigned long long)(line + 1) << 8) | (unsigned)code);
}

// ---- Criteo (learn/base/criteo_parser.h:47-87), one thread per line ----------------------------------------
// A block owns kCriteoLPB consecutive lines = one contiguous span of text, which it first copies to shared memory
// with coalesced 16-byte loads (a thread walking its line in global memory would touch a different 128-byte line
// of the cache than each of its 31 neighbours on every byte). Spans beyond the buffer are parsed in place.
constexpr int kCriteoLPB = 128;
constexpr int kCriteoSpan = 56 * 1024;   // bytes of text staged per block (average Criteo line: ~230 bytes)
__global__ void __launch_bounds__(kCriteoLPB) criteo_lines_kernel(const char* text, size_t len, const uint32_t* starts,
                                                                  size_t lines, int is_train, float* label, unsigned* cnt,
                                                                  uint64_t* tmp, unsigned long long* err) {
  extern __shared__ uint4 stage[];
  const size_t l0 = (size_t)blockIdx.x * kCriteoLPB;
  const size_t r = l0 + threadIdx.x;
  const size_t g0 = starts[l0], g1 = l0 + kCriteoLPB < lines ? starts[l0 + kCriteoLPB] : len;
  const size_t a0 = g0 & ~(size_t)15;                       // 16-byte aligned start of the copy
  const char* base = text;                                  // address of text byte 0 in the space we parse from
  const bool aligned = ((uintptr_t)text & 15) == 0;
  if (aligned && g1 - a0 <= (size_t)kCriteoSpan) {
    const size_t n16 = (g1 - a0 + 15) / 16;
    const uint4* src = reinterpret_cast<const uint4*>(text + a0);
    for (size_t k = threadIdx.x; k < n16; k += kCriteoLPB) {
      if (a0 + 16 * k + 16 <= len) stage[k] = __ldg(src + k);
      else for (size_t b = a0 + 16 * k; b < len; ++b) reinterpret_cast<char*>(stage)[b - a0] = text[b];   // the text's last bytes
    }
    __syncthreads();
    base = reinterpret_cast<const char*>(stage) - a0;
  }
  if (r >= lines) return;
  const char *p = base + starts[r], *end = base + len;
  const char* lend = line_end(p, end);
  uint64_t* out = tmp + r * kCriteoMax;
  unsigned n = 0;
  int bad = 0;
  cnt[r] = 0;
  if (is_train) {                                                       // :54-62
    const char* pp = p;
    while (pp != lend && *pp != '\t') ++pp;
    if (pp == lend || pp == p) { report(err, 1, r); return; }
    label[r] = (float)parse_number(p, pp, &bad);
    if (bad) { report(err, 9, r); return; }
    p = pp + 1;
  } else {
    label[r] = 0.f;
  }
  for (int i = 0; i < kCriteoInts; ++i) {                               // :65-72
    const char* pp = p;
    while (pp != lend && *pp != '\t') ++pp;
    if (pp == lend) { report(err, 2, r); return; }
    if (pp > p) {
      if (pp - p > 16) { report(err, 3, r); return; }
      out[n++] = (cityhash64_short(p, (int)(pp - p)) >> 10) | ((uint64_t)i << 54);
    }
    p = pp + 1;
  }
  int i = 0;
  for (; i < kCriteoCats; ++i) {                                        // :75-85
    if (p == end) break;
    if (p == lend) { if (i != kCriteoCats - 1) { report(err, 4, r); return; } break; }
    if (*p == '\t') { ++p; continue; }
    if (*p == ' ' || *p == '\v' || *p == '\f') { report(err, 5, r); return; }
    const char* pp = p + 8;
    if (pp > lend) { report(err, 6, r); return; }
    if (pp != end && !(*pp == '\t' || is_brk(*pp) || *pp == ' ' || *pp == '\v' || *pp == '\f')) { report(err, 6, r); return; }
    out[n++] = (cityhash64_short(p, 8) >> 10) | ((uint64_t)(i + kCriteoInts) << 54);
    p = pp == end ? end : pp + 1;
    if (pp == end || is_brk(*pp)) break;
  }
  if (i == kCriteoCats && p < lend) { report(err, 7, r); return; }
  cnt[r] = n;
}
struct CntFlag {
  const unsigned* cnt;
  __device__ unsigned operator()(size_t i) const { return cnt[i]; }
};
struct EmitOffset {
  size_t* offset;
  __device__ void operator()(size_t r, unsigned v, unsigned prefix) const { offset[r + 1] = (size_t)prefix + v; }
};
// one warp moves one line's strip (<= 39 ids, contiguous on both sides) to its place in the CSR
__global__ void __launch_bounds__(256) compact_criteo_kernel(const uint64_t* tmp, const size_t* offset, size_t lines,
                                                             uint64_t* index) {
  const size_t r = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned lane = threadIdx.x & 31;
  if (r >= lines) return;
  const size_t o = offset[r], n = offset[r + 1] - o;
  const uint64_t* s = tmp + r * kCriteoMax;
  for (size_t k = lane; k < n; k += 32) index[o + k] = s[k];
}

// ---- libsvm, one thread per line ---------------------------------------------------------------------------
// value of the scan: (1 << 32 | tokens) for a line that holds a row, 0 for a blank one
struct LibsvmCount {
  const char* text;
  size_t len;
  const uint32_t* starts;
  __device__ unsigned long long operator()(size_t r) const {
    const char *p = text + starts[r], *end = text + len;
    const char* lend = line_end(p, end);
    while (p != lend && is_blank(*p)) ++p;
    if (p == lend) return 0ULL;
    unsigned tok = 0;
    bool in = false;
    for (; p != lend; ++p) { const bool b = is_blank(*p); if (!b && !in) ++tok; in = !b; }
    return (1ULL << 32) | (tok - 1);
  }
};
struct LibsvmEmit {
  const char* text;
  size_t len;
  const uint32_t* starts;
  float* label;
  size_t* offset;
  uint64_t* index;
  float* value;
  unsigned long long* nvalue;
  unsigned long long* err;
  __device__ void operator()(size_t line, unsigned long long v, unsigned long long prefix) const {
    if (!v) return;
    const size_t r = (size_t)(prefix >> 32);
    size_t o = (size_t)(prefix & 0xffffffffULL);
    offset[r + 1] = o + (size_t)(v & 0xffffffffULL);
    const char *p = text + starts[line], *end = text + len;
    const char* lend = line_end(p, end);
    int bad = 0;
    unsigned nv = 0;
    while (p != lend && is_blank(*p)) ++p;
    const char* q = p;
    while (q != lend && !is_blank(*q)) ++q;
    label[r] = (float)parse_number(p, q, &bad);
    p = q;
    for (;;) {
      while (p != lend && is_blank(*p)) ++p;
      if (p == lend) break;
      q = p;
      const char* c = nullptr;
      while (q != lend && !is_blank(*q)) { if (*q == ':' && !c) c = q; ++q; }
      index[o] = parse_u64(p, c ? c : q);
      if (c) { value[o] = (float)parse_number(c + 1, q, &bad); ++nv; }
      ++o;
      p = q;
    }
    if (bad) report(err, 9, line);
    if (nv) atomicAdd(nvalue, (unsigned long long)nv);
  }
};

// ---- adfea (learn/base/adfea_parser.h:38-82), one thread per line ------------------------------------------
// The reference walks a token stream: a token "idx:gid" is a feature, index = idx >> 10 | gid << 54; every third
// token without a colon opens a row, its first character == '1' being the label (the two before it are the line id
// and a count). With one row per line that is: three plain tokens, then features. Lines of any other shape, and
// tokens that do not start with a digit (CHECK_NE(head, p)), fail the call.
__device__ __forceinline__ bool is_space_adfea(char c) { return c == ' ' || c == '\t' || c == '\v' || c == '\f'; }
struct AdfeaCount {
  const char* text;
  size_t len;
  const uint32_t* starts;
  __device__ unsigned long long operator()(size_t r) const {
    const char *p = text + starts[r], *end = text + len;
    const char* lend = line_end(p, end);
    unsigned tok = 0;
    bool in = false;
    for (; p != lend; ++p) { const bool b = is_space_adfea(*p); if (!b && !in) ++tok; in = !b; }
    if (tok == 0) return 0ULL;                          // blank line
    return (1ULL << 32) | (tok > 3 ? tok - 3 : 0u);
  }
};
struct AdfeaEmit {
  const char* text;
  size_t len;
  const uint32_t* starts;
  float* label;
  size_t* offset;
  uint64_t* index;
  unsigned long long* err;
  __device__ void operator()(size_t line, unsigned long long v, unsigned long long prefix) const {
    if (!v) return;
    const size_t r = (size_t)(prefix >> 32);
    size_t o = (size_t)(prefix & 0xffffffffULL);
    offset[r + 1] = o + (size_t)(v & 0xffffffffULL);
    const char *p = text + starts[line], *end = text + len;
    const char* lend = line_end(p, end);
    unsigned t = 0;
    bool bad = false;
    label[r] = 0.f;
    for (;;) {
      while (p != lend && is_space_adfea(*p)) ++p;
      if (p == lend) break;
      const char* head = p;
      while (p != lend && is_digit(*p)) ++p;
      if (p == head) { bad = true; break; }                                   // CHECK_NE(head, p)
      if (p != lend && *p == ':') {
        const char* g = ++p;
        while (p != lend && is_digit(*p)) ++p;
        if (t < 3 || (p != lend && !is_space_adfea(*p))) { bad = true; break; }
        index[o++] = (parse_u64(head, g - 1) >> 10) | (parse_u64(g, p) << 54);  // :52-56
      } else {
        if (t >= 3 || (p != lend && !is_space_adfea(*p))) { bad = true; break; }
        if (t == 2) label[r] = *head == '1' ? 1.f : 0.f;                        // :65
      }
      ++t;
    }
    if (bad || t < 3) report(err, t < 3 && !bad ? 11 : 10, line);
  }
};

// ---- container kernels -------------------------------------------------------------------------------------
__global__ void rebase_offsets_kernel(const size_t* src, size_t n, size_t src_base, size_t dst_base, size_t* dst) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[i] - src_base + dst_base;
}
__global__ void fill_f32_kernel(float* p, size_t n, float v) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
// one warp copies one listed row
__global__ void gather_rows_kernel(const uint32_t* ids, const size_t* dst_off, size_t n, const size_t* s_off,
                                   const float* s_label, const uint64_t* s_index, const float* s_value,
                                   size_t* d_off, float* d_label, uint64_t* d_index, float* d_value, size_t row0) {
  const size_t w = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned lane = threadIdx.x & 31;
  if (w >= n) return;
  const size_t r = ids[w], sb = s_off[r], m = s_off[r + 1] - sb, db = dst_off[w];
  if (lane == 0) { d_off[row0 + w + 1] = db + m; d_label[row0 + w] = s_label[r]; }
  for (size_t k = lane; k < m; k += 32) {
    d_index[db + k] = s_index[sb + k];
    if (d_value) d_value[db + k] = s_value ? s_value[sb + k] : 1.f;
  }
}
__global__ void any_not_one_kernel(const float* v, size_t n, unsigned* flag) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool x = i < n && v[i] != 1.f;
  if (__any_sync(0xffffffffu, x) && (threadIdx.x & 31) == 0) *flag = 1u;
}

}  // namespace
}  // namespace wh

using namespace wh;

namespace {

int rows_reserve(whRows* r, size_t rows, size_t nnz, bool value, bool keep) {
  // grow-only buffers; `keep` preserves the current contents across a reallocation
  auto grow = [&](DevBuf& b, size_t bytes, size_t used) -> int {
    if (bytes <= b.cap) return WH_OK;
    if (!keep || used == 0) return b.reserve(bytes);
    DevBuf n;
    WH_TRY(n.reserve(bytes + bytes / 2));
    WH_CUDA(cudaMemcpyAsync(n.p, b.p, used, cudaMemcpyDeviceToDevice, r->ctx->stream));
    WH_CUDA(cudaStreamSynchronize(r->ctx->stream));
    b.release();
    b = n;
    return WH_OK;
  };
  WH_TRY(grow(r->offset, (rows + 1) * sizeof(size_t), (r->rows + 1) * sizeof(size_t)));
  WH_TRY(grow(r->label, (rows ? rows : 1) * sizeof(float), r->rows * sizeof(float)));
  WH_TRY(grow(r->index, (nnz ? nnz : 1) * sizeof(uint64_t), r->nnz * sizeof(uint64_t)));
  if (value) WH_TRY(grow(r->value, (nnz ? nnz : 1) * sizeof(float), r->has_value ? r->nnz * sizeof(float) : 0));
  return WH_OK;
}

int host_offsets(const whRows* cr) {
  whRows* r = const_cast<whRows*>(cr);
  if (r->h_valid) return WH_OK;
  r->h_offset.resize(r->rows + 1);
  WH_CUDA(cudaMemcpyAsync(r->h_offset.data(), r->offset.p, (r->rows + 1) * sizeof(size_t), cudaMemcpyDeviceToHost,
                          r->ctx->stream));
  WH_CUDA(cudaStreamSynchronize(r->ctx->stream));
  r->h_valid = true;
  return WH_OK;
}

const char* parse_error_text(int code) {
  switch (code) {
    case 1: return "no label field (criteo_parser.h:57 CHECK_NE; a test file needs format criteo_test)";
    case 2: return "fewer than 13 integer fields on the line";
    case 3: return "integer field longer than 16 bytes";
    case 4: return "line ends in an empty categorical field before the 26th";
    case 5: return "categorical field starts with white space";
    case 6: return "categorical field is not 8 bytes followed by white space (criteo_parser.h:79 CHECK)";
    case 7: return "bytes after the 26th categorical field";
    case 8: return "some features carry a value and some do not";
    case 9: return "inf / nan / hexadecimal numeric literal";
    case 10: return "adfea: a token is neither digits nor digits:digits, or features and plain tokens are out of order";
    case 11: return "adfea: fewer than three leading tokens (line id, count, label)";
  }
  return "malformed input";
}

}  // namespace

namespace wh {
// RowBlockContainer::Push(RowBlock) from host arrays: `rows` rows appended to a device container. offset is relative
// to offset[0]; value may be NULL (a mix of blocks with and without values is completed with ones, as wh_rows_append).
int rows_append_host(whContext* ctx, whRows* dst, size_t rows, const size_t* offset, const float* label, const uint64_t* index,
                     const float* value) {
  cudaStream_t st = ctx->stream;
  if (rows == 0) return WH_OK;
  const size_t add = offset[rows] - offset[0];
  const bool was_empty = dst->rows == 0;
  const bool want_value = was_empty ? value != nullptr : (dst->has_value || value != nullptr);
  WH_TRY(rows_reserve(dst, dst->rows + rows, dst->nnz + add, want_value, true));
  if (want_value && !dst->has_value && dst->nnz)
    WH_LAUNCH(fill_f32_kernel, (unsigned)div_up(dst->nnz, 256), 256, 0, st, dst->value.as<float>(), dst->nnz, 1.f);
  dst->has_value = want_value;
  std::vector<size_t> off(rows + (was_empty ? 1 : 0));
  size_t* o = off.data();
  if (was_empty) *o++ = 0;
  for (size_t i = 0; i < rows; ++i) o[i] = dst->nnz + (offset[i + 1] - offset[0]);
  WH_CUDA(cudaMemcpyAsync(dst->offset.as<size_t>() + (was_empty ? 0 : dst->rows + 1), off.data(), off.size() * sizeof(size_t),
                          cudaMemcpyHostToDevice, st));
  WH_CUDA(cudaMemcpyAsync(dst->label.as<float>() + dst->rows, label, rows * sizeof(float), cudaMemcpyHostToDevice, st));
  if (add) {
    WH_CUDA(cudaMemcpyAsync(dst->index.as<uint64_t>() + dst->nnz, index + offset[0], add * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
    if (want_value && value) WH_CUDA(cudaMemcpyAsync(dst->value.as<float>() + dst->nnz, value + offset[0], add * sizeof(float), cudaMemcpyHostToDevice, st));
    if (want_value && !value) WH_LAUNCH(fill_f32_kernel, (unsigned)div_up(add, 256), 256, 0, st, dst->value.as<float>() + dst->nnz, add, 1.f);
  }
  WH_CUDA(cudaStreamSynchronize(st));   // the host arrays are the caller's temporaries
  dst->rows += rows; dst->nnz += add;
  dst->h_valid = false;
  return WH_OK;
}
}  // namespace wh

extern "C" {

int wh_rows_create(whContext* ctx, whRows** out) {
  WH_REQUIRE(ctx && out, "wh_rows_create: null argument");
  DeviceGuard g(ctx->device);
  whRows* r = new whRows;
  r->ctx = ctx;
  int rc = r->offset.reserve(sizeof(size_t));
  if (rc == WH_OK && cudaMemsetAsync(r->offset.p, 0, sizeof(size_t), ctx->stream) != cudaSuccess) rc = WH_ERR_CUDA;
  if (rc != WH_OK) { delete r; return rc; }
  *out = r;
  return WH_OK;
}

void wh_rows_destroy(whRows* r) {
  if (!r) return;
  DeviceGuard g(r->ctx->device);
  cudaStreamSynchronize(r->ctx->stream);
  delete r;
}

int wh_rows_clear(whRows* r) {
  WH_REQUIRE(r, "wh_rows_clear: null argument");
  DeviceGuard g(r->ctx->device);
  r->rows = r->nnz = 0;
  r->has_value = false;
  r->h_valid = false;
  WH_CUDA(cudaMemsetAsync(r->offset.p, 0, sizeof(size_t), r->ctx->stream));
  return WH_OK;
}

size_t wh_rows_size(const whRows* r) { return r ? r->rows : 0; }
size_t wh_rows_nnz(const whRows* r) { return r ? r->nnz : 0; }
int wh_rows_has_value(const whRows* r) { return r && r->has_value; }

int wh_rows_view(const whRows* r, whRowBlock* view) {
  WH_REQUIRE(r && view, "wh_rows_view: null argument");
  view->size = r->rows;
  view->offset = r->offset.as<size_t>();
  view->label = r->label.as<float>();
  view->weight = nullptr;
  view->index = r->index.as<uint64_t>();
  view->value = r->has_value ? r->value.as<float>() : nullptr;
  return WH_OK;
}

int wh_rows_copy(whContext* ctx, const whRows* r, size_t* offset, float* label, uint64_t* index, float* value) {
  WH_REQUIRE(ctx && r, "wh_rows_copy: null argument");
  DeviceGuard g(ctx->device);
  cudaStream_t s = r->ctx->stream;
  if (offset) WH_CUDA(cudaMemcpyAsync(offset, r->offset.p, (r->rows + 1) * sizeof(size_t), cudaMemcpyDeviceToHost, s));
  if (label && r->rows) WH_CUDA(cudaMemcpyAsync(label, r->label.p, r->rows * sizeof(float), cudaMemcpyDeviceToHost, s));
  if (index && r->nnz) WH_CUDA(cudaMemcpyAsync(index, r->index.p, r->nnz * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
  if (value && r->nnz && r->has_value)
    WH_CUDA(cudaMemcpyAsync(value, r->value.p, r->nnz * sizeof(float), cudaMemcpyDeviceToHost, s));
  WH_CUDA(cudaStreamSynchronize(s));
  return WH_OK;
}

int wh_parse_text(whContext* ctx, int format, const char* text, size_t len, int mem, whRows** out) {
  WH_REQUIRE(ctx && out && (text || len == 0), "wh_parse_text: null argument");
  WH_REQUIRE(format == WH_FMT_LIBSVM || format == WH_FMT_CRITEO || format == WH_FMT_CRITEO_TEST || format == WH_FMT_ADFEA,
             "wh_parse_text: unknown format %d (crb records go through wh_parse_crb)", format);
  WH_REQUIRE(mem == WH_HOST || mem == WH_DEVICE, "wh_parse_text: bad mem %d", mem);
  WH_REQUIRE(len < (1ULL << 32), "wh_parse_text: a chunk is at most 4 GiB - 1 bytes (got %zu)", len);
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  // *out == NULL: a new container; otherwise the caller's container is cleared and refilled (its buffers are reused)
  whRows* r = *out;
  const bool fresh = r == nullptr;
  if (fresh) WH_TRY(wh_rows_create(ctx, &r)); else WH_TRY(wh_rows_clear(r));
  struct Guard { whRows* r; ~Guard() { if (r) wh_rows_destroy(r); } } guard{fresh ? r : nullptr};
  if (len == 0) { guard.r = nullptr; *out = r; return WH_OK; }

  const char* d_text = nullptr;
  WH_TRY(to_device(ctx, ctx->stage[0], text, len, mem, &d_text));
  // device scalars: [0] lines (u32), [1] reserved, then u64: [1] packed totals, [2] nvalue, [3] err
  WH_TRY(ctx->small.reserve(64));
  unsigned* d_lines = ctx->small.as<unsigned>();
  unsigned long long* d_u64 = ctx->small.as<unsigned long long>();
  WH_CUDA(cudaMemsetAsync(d_u64, 0, 24, st));
  WH_CUDA(cudaMemsetAsync(d_u64 + 3, 0xff, 8, st));

  // 1. line starts: count, then emit
  const unsigned nb = (unsigned)div_up(len, kScanTile);
  WH_TRY(ctx->scan_tmp.reserve((size_t)nb * sizeof(unsigned long long)));
  unsigned* sums = ctx->scan_tmp.as<unsigned>();
  LineStartFlag lf{d_text};
  WH_LAUNCH((scan_tile_sums_kernel<unsigned, LineStartFlag>), nb, kScanThreads, 0, st, lf, len, sums);
  WH_LAUNCH((scan_sums_kernel<unsigned>), 1, 1024, 0, st, sums, (size_t)nb, d_lines);
  unsigned* h = ctx->pinned.as<unsigned>();
  WH_CUDA(cudaMemcpyAsync(h, d_lines, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
  WH_CUDA(cudaStreamSynchronize(st));
  const size_t lines = h[0];
  if (lines == 0) { guard.r = nullptr; *out = r; return WH_OK; }
  WH_TRY(ctx->stage[1].reserve(lines * sizeof(uint32_t)));
  uint32_t* starts = ctx->stage[1].as<uint32_t>();
  WH_LAUNCH((scan_apply_kernel<unsigned, LineStartFlag, EmitLineStart>), nb, kScanThreads, 0, st, lf,
            EmitLineStart{starts}, len, sums);

  unsigned long long* hp = ctx->pinned.as<unsigned long long>();
  if (format == WH_FMT_ADFEA) {
    // as libsvm below, with the adfea functors and no values
    const unsigned nbl = (unsigned)div_up(lines, kScanTile);
    WH_TRY(ctx->stage[2].reserve((size_t)nbl * sizeof(unsigned long long)));
    unsigned long long* sums2 = ctx->stage[2].as<unsigned long long>();
    AdfeaCount cf{d_text, len, starts};
    WH_LAUNCH((scan_tile_sums_kernel<unsigned long long, AdfeaCount>), nbl, kScanThreads, 0, st, cf, lines, sums2);
    WH_LAUNCH((scan_sums_kernel<unsigned long long>), 1, 1024, 0, st, sums2, (size_t)nbl, d_u64 + 1);
    WH_CUDA(cudaMemcpyAsync(hp, d_u64 + 1, 8, cudaMemcpyDeviceToHost, st));
    WH_CUDA(cudaStreamSynchronize(st));
    const size_t rows = (size_t)(hp[0] >> 32), nnz = (size_t)(hp[0] & 0xffffffffULL);
    WH_TRY(rows_reserve(r, rows, nnz, false, false));
    WH_CUDA(cudaMemsetAsync(r->offset.p, 0, sizeof(size_t), st));
    AdfeaEmit em{d_text, len, starts, r->label.as<float>(), r->offset.as<size_t>(), r->index.as<uint64_t>(), d_u64 + 3};
    WH_LAUNCH((scan_apply_kernel<unsigned long long, AdfeaCount, AdfeaEmit>), nbl, kScanThreads, 0, st, cf, em, lines, sums2);
    WH_CUDA(cudaMemcpyAsync(hp + 1, d_u64 + 3, 8, cudaMemcpyDeviceToHost, st));
    WH_CUDA(cudaStreamSynchronize(st));
    r->rows = rows;
    r->nnz = nnz;
    r->has_value = false;
  } else if (format == WH_FMT_LIBSVM) {
    // 2. tokens per line -> row number and offset of every line; 3. parse into place
    const unsigned nbl = (unsigned)div_up(lines, kScanTile);
    WH_TRY(ctx->stage[2].reserve((size_t)nbl * sizeof(unsigned long long)));
    unsigned long long* sums2 = ctx->stage[2].as<unsigned long long>();
    LibsvmCount cf{d_text, len, starts};
    WH_LAUNCH((scan_tile_sums_kernel<unsigned long long, LibsvmCount>), nbl, kScanThreads, 0, st, cf, lines, sums2);
    WH_LAUNCH((scan_sums_kernel<unsigned long long>), 1, 1024, 0, st, sums2, (size_t)nbl, d_u64 + 1);
    WH_CUDA(cudaMemcpyAsync(hp, d_u64 + 1, 8, cudaMemcpyDeviceToHost, st));
    WH_CUDA(cudaStreamSynchronize(st));
    const size_t rows = (size_t)(hp[0] >> 32), nnz = (size_t)(hp[0] & 0xffffffffULL);
    WH_TRY(rows_reserve(r, rows, nnz, true, false));
    WH_CUDA(cudaMemsetAsync(r->offset.p, 0, sizeof(size_t), st));
    LibsvmEmit em{d_text, len, starts, r->label.as<float>(), r->offset.as<size_t>(), r->index.as<uint64_t>(),
                  r->value.as<float>(), d_u64 + 2, d_u64 + 3};
    WH_LAUNCH((scan_apply_kernel<unsigned long long, LibsvmCount, LibsvmEmit>), nbl, kScanThreads, 0, st, cf, em,
              lines, sums2);
    WH_CUDA(cudaMemcpyAsync(hp, d_u64 + 2, 16, cudaMemcpyDeviceToHost, st));
    WH_CUDA(cudaStreamSynchronize(st));
    r->rows = rows;
    r->nnz = nnz;
    if (hp[1] == ~0ULL && !(hp[0] == 0 || hp[0] == nnz)) hp[1] = (1ULL << 8) | 8;
    r->has_value = nnz != 0 && hp[0] == nnz;
  } else {
    // 2. parse every line into a 39-slot strip; 3. compact the strips by a scan over the row lengths
    WH_TRY(rows_reserve(r, lines, lines * kCriteoMax, false, false));
    WH_CUDA(cudaMemsetAsync(r->offset.p, 0, sizeof(size_t), st));
    WH_TRY(ctx->stage[2].reserve(lines * kCriteoMax * sizeof(uint64_t)));
    WH_TRY(ctx->stage[3].reserve(lines * sizeof(unsigned)));
    const unsigned nbl = (unsigned)div_up(lines, kScanTile);
    WH_TRY(ctx->stage[4].reserve((size_t)nbl * sizeof(unsigned)));
    unsigned* cnt = ctx->stage[3].as<unsigned>();
    uint64_t* tmp = ctx->stage[2].as<uint64_t>();
    // > 48 KiB of dynamic shared memory needs the opt-in (per device, so repeated; it is a host-side table write)
    WH_CUDA(cudaFuncSetAttribute(criteo_lines_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kCriteoSpan + 16));
    WH_LAUNCH(criteo_lines_kernel, (unsigned)div_up(lines, kCriteoLPB), kCriteoLPB, kCriteoSpan + 16, st, d_text, len, starts, lines,
              format == WH_FMT_CRITEO ? 1 : 0, r->label.as<float>(), cnt, tmp, d_u64 + 3);
    unsigned* d_nnz = ctx->small.as<unsigned>() + 1;
    WH_TRY((scan_and_apply<unsigned, CntFlag, EmitOffset>(ctx, CntFlag{cnt}, EmitOffset{r->offset.as<size_t>()}, lines,
                                                          ctx->stage[4].as<unsigned>(), d_nnz)));
    WH_LAUNCH(compact_criteo_kernel, (unsigned)div_up(lines * 32, 256), 256, 0, st, tmp, r->offset.as<size_t>(), lines,
              r->index.as<uint64_t>());
    WH_CUDA(cudaMemcpyAsync(hp, d_u64 + 3, 8, cudaMemcpyDeviceToHost, st));
    WH_CUDA(cudaMemcpyAsync(hp + 1, d_nnz, 4, cudaMemcpyDeviceToHost, st));
    WH_CUDA(cudaStreamSynchronize(st));
    r->rows = lines;
    r->nnz = ((unsigned*)(hp + 1))[0];
    r->has_value = false;
    hp[1] = hp[0];
  }
  if (hp[1] != ~0ULL) {
    set_error("wh_parse_text: line %llu: %s", (unsigned long long)(hp[1] >> 8), parse_error_text((int)(hp[1] & 0xff)));
    return WH_ERR_INVALID;
  }
  guard.r = nullptr;
  *out = r;
  return WH_OK;
}

int wh_rows_append(whContext* ctx, whRows* dst, const whRows* src, size_t begin, size_t n, const uint32_t* row_ids) {
  WH_REQUIRE(ctx && dst && src && dst != src, "wh_rows_append: null or aliased argument");
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  if (n == 0) return WH_OK;
  WH_TRY(host_offsets(src));
  const std::vector<size_t>& so = src->h_offset;
  size_t add = 0;
  if (row_ids) {
    for (size_t i = 0; i < n; ++i) {
      WH_REQUIRE(row_ids[i] < src->rows, "wh_rows_append: row %u of %zu", row_ids[i], src->rows);
      add += so[row_ids[i] + 1] - so[row_ids[i]];
    }
  } else {
    WH_REQUIRE(begin + n <= src->rows, "wh_rows_append: rows [%zu, %zu) of %zu", begin, begin + n, src->rows);
    add = so[begin + n] - so[begin];
  }
  // values: an empty container takes the source's kind; a mix is completed with ones
  const bool was_empty = dst->rows == 0;
  const bool want_value = was_empty ? src->has_value : (dst->has_value || src->has_value);
  WH_TRY(rows_reserve(dst, dst->rows + n, dst->nnz + add, want_value, true));
  if (want_value && !dst->has_value && dst->nnz)
    WH_LAUNCH(fill_f32_kernel, (unsigned)div_up(dst->nnz, 256), 256, 0, st, dst->value.as<float>(), dst->nnz, 1.f);
  dst->has_value = want_value;
  float* dv = want_value ? dst->value.as<float>() : nullptr;
  const float* sv = src->has_value ? src->value.as<float>() : nullptr;
  if (row_ids) {
    // host: destination offsets of the listed rows; device: one warp per row
    std::vector<size_t> doff(n);
    size_t o = dst->nnz;
    for (size_t i = 0; i < n; ++i) { doff[i] = o; o += so[row_ids[i] + 1] - so[row_ids[i]]; }
    WH_TRY(ctx->stage[5].reserve(n * sizeof(uint32_t)));
    WH_TRY(ctx->stage[6].reserve(n * sizeof(size_t)));
    WH_CUDA(cudaMemcpyAsync(ctx->stage[5].p, row_ids, n * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    WH_CUDA(cudaMemcpyAsync(ctx->stage[6].p, doff.data(), n * sizeof(size_t), cudaMemcpyHostToDevice, st));
    WH_LAUNCH(gather_rows_kernel, (unsigned)div_up(n * 32, 256), 256, 0, st, ctx->stage[5].as<uint32_t>(),
              ctx->stage[6].as<size_t>(), n, src->offset.as<size_t>(), src->label.as<float>(),
              src->index.as<uint64_t>(), sv, dst->offset.as<size_t>(), dst->label.as<float>(),
              dst->index.as<uint64_t>(), dv, dst->rows);
    WH_CUDA(cudaStreamSynchronize(st));   // doff / row_ids are pageable host memory
  } else {
    const size_t sb = so[begin];
    WH_LAUNCH(rebase_offsets_kernel, (unsigned)div_up(n, 256), 256, 0, st, src->offset.as<size_t>() + begin + 1, n, sb,
              dst->nnz, dst->offset.as<size_t>() + dst->rows + 1);
    WH_CUDA(cudaMemcpyAsync(dst->label.as<float>() + dst->rows, src->label.as<float>() + begin, n * sizeof(float),
                            cudaMemcpyDeviceToDevice, st));
    if (add) {
      WH_CUDA(cudaMemcpyAsync(dst->index.as<uint64_t>() + dst->nnz, src->index.as<uint64_t>() + sb,
                              add * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
      if (dv && sv) WH_CUDA(cudaMemcpyAsync(dv + dst->nnz, sv + sb, add * sizeof(float), cudaMemcpyDeviceToDevice, st));
      if (dv && !sv) WH_LAUNCH(fill_f32_kernel, (unsigned)div_up(add, 256), 256, 0, st, dv + dst->nnz, add, 1.f);
    }
  }
  dst->rows += n;
  dst->nnz += add;
  dst->h_valid = false;
  return WH_OK;
}

int wh_rows_drop_unit_values(whContext* ctx, whRows* r) {
  WH_REQUIRE(ctx && r, "wh_rows_drop_unit_values: null argument");
  if (!r->has_value) return WH_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  WH_TRY(ctx->small.reserve(64));
  unsigned* flag = ctx->small.as<unsigned>();
  WH_CUDA(cudaMemsetAsync(flag, 0, 4, st));
  WH_LAUNCH(any_not_one_kernel, (unsigned)div_up(r->nnz, 256), 256, 0, st, r->value.as<float>(), r->nnz, flag);
  unsigned* h = ctx->pinned.as<unsigned>();
  WH_CUDA(cudaMemcpyAsync(h, flag, 4, cudaMemcpyDeviceToHost, st));
  WH_CUDA(cudaStreamSynchronize(st));
  if (!h[0]) r->has_value = false;      // learn/base/minibatch_iter.h:116-118
  return WH_OK;
}

}  // extern "C"
