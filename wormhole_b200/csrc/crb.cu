// The `crb` input format (learn/base/crb_parser.h:27-40, learn/base/compressed_row_block.h:44-69,116-137): RecordIO
// records, each one a CompressedRowBlock = {magic 1196140743, sizeof(IndexType), nrows} followed by five arrays
// (label f32[nrows], offset size_t[nrows+1], index IndexType[nnz], value f32[nnz], weight f32[nrows]), each stored as
// {int compressed_size, LZ4 block}; compressed_size <= 0 means "absent".
//
// Two pieces of that path live in dependencies the reference does not vendor, restated here from their published
// formats: LZ4's block format (lz4 r131, make/deps.mk) and dmlc-core's RecordIO framing (dmlc/recordio.h: records
// start with the magic 0xced7230a and a word {cflag : 3 | length : 29}, are padded to 4 bytes, and a record whose
// payload contains the magic at a 4-byte boundary is split there into parts with cflag 1 (first), 2 (middle), 3
// (last), the magic being the part separator). tests/test_crb.py pins the LZ4 decoder against liblz4 itself (blocks
// compressed by pyarrow's lz4_raw codec).
//
// An LZ4 block is a strictly sequential byte code -- every token depends on where the previous one ended -- and a
// record holds five of them, so the decode runs on the host, straight from the file buffer; the decoded arrays cross
// PCIe once, into the device container that MinibatchIter assembles minibatches from (the text formats go the other
// way round: text crosses, parsed in HBM).
#include <string.h>

#include <vector>

#include "common.cuh"

struct whRows;
extern "C" {
int wh_rows_create(whContext* ctx, whRows** out);
void wh_rows_destroy(whRows* r);
int wh_rows_clear(whRows* r);
}
namespace wh {
// appends host arrays (rows, offsets relative to offset[0]) to a device container (parse.cu)
int rows_append_host(whContext* ctx, whRows* r, size_t rows, const size_t* offset, const float* label, const uint64_t* index,
                     const float* value);

// LZ4_decompress_safe: decodes exactly one block into dst[0, cap); returns the decoded size or -1 on malformed input.
// Block format: sequences of {token: literal length (high nibble) | match length - 4 (low nibble)}, each nibble
// extended by 255-valued bytes when it is 15; literals; a 2-byte little-endian match offset; the last sequence ends
// after its literals.
long lz4_block_decode(const unsigned char* src, size_t n, unsigned char* dst, size_t cap) {
  size_t ip = 0, op = 0;
  while (ip < n) {
    const unsigned token = src[ip++];
    size_t lit = token >> 4;
    if (lit == 15) {
      unsigned b;
      do { if (ip >= n) return -1; b = src[ip++]; lit += b; } while (b == 255);
    }
    if (lit > n - ip || lit > cap - op) return -1;
    memcpy(dst + op, src + ip, lit);
    ip += lit; op += lit;
    if (ip >= n) break;                          // the last sequence has no match
    if (n - ip < 2) return -1;
    const size_t off = (size_t)src[ip] | ((size_t)src[ip + 1] << 8);
    ip += 2;
    if (off == 0 || off > op) return -1;
    size_t len = token & 15;
    if (len == 15) {
      unsigned b;
      do { if (ip >= n) return -1; b = src[ip++]; len += b; } while (b == 255);
    }
    len += 4;
    if (len > cap - op) return -1;
    for (size_t i = 0; i < len; ++i) dst[op + i] = dst[op + i - off];   // may overlap: byte by byte
    op += len;
  }
  return (long)op;
}

constexpr uint32_t kRecordMagic = 0xced7230au;
constexpr int kCrbMagic = 1196140743;

// One CompressedRowBlock record -> appended to `out`. Follows Decompress, compressed_row_block.h:51-69,116-126.
static int crb_record(whContext* ctx, const unsigned char* p, size_t n, whRows* out, std::vector<unsigned char>& scratch) {
  size_t cur = 0;
  auto read_int = [&](int* v) { if (cur + 4 > n) return false; memcpy(v, p + cur, 4); cur += 4; return true; };
  int magic, isz, nrows;
  WH_REQUIRE(read_int(&magic) && magic == kCrbMagic, "crb: wrong data format (magic %d)", magic);
  WH_REQUIRE(read_int(&isz) && isz == 8, "crb: wrong indextype (%d bytes; this reader takes uint64 ids)", isz);
  WH_REQUIRE(read_int(&nrows) && nrows >= 0, "crb: bad row count");
  // label | offset | index | value | weight
  std::vector<float> label, value;
  std::vector<size_t> offset;
  std::vector<uint64_t> index;
  auto array = [&](void* dst, size_t bytes, bool* present) -> int {
    int cp;
    WH_REQUIRE(read_int(&cp), "crb: truncated record");
    *present = cp > 0;
    if (cp <= 0) return WH_OK;
    WH_REQUIRE(cur + (size_t)cp <= n, "crb: an array runs past its record");
    const long got = lz4_block_decode(p + cur, (size_t)cp, static_cast<unsigned char*>(dst), bytes);
    WH_REQUIRE(got == (long)bytes, "crb: LZ4 block decodes to %ld bytes, expected %zu", got, bytes);
    cur += (size_t)cp;
    return WH_OK;
  };
  bool has;
  label.resize((size_t)nrows);
  WH_TRY(array(label.data(), (size_t)nrows * 4, &has));
  if (!has) label.assign((size_t)nrows, 0.f);
  offset.resize((size_t)nrows + 1);
  WH_TRY(array(offset.data(), ((size_t)nrows + 1) * sizeof(size_t), &has));
  WH_REQUIRE(has, "crb: a record without offsets");
  for (int i = 0; i < nrows; ++i) WH_REQUIRE(offset[i] <= offset[i + 1], "crb: offsets decrease");
  const size_t nnz = offset[nrows] - offset[0];
  index.resize(nnz);
  WH_TRY(array(index.data(), nnz * 8, &has));
  WH_REQUIRE(has || nnz == 0, "crb: a record without indices");
  value.resize(nnz);
  bool has_value;
  WH_TRY(array(value.data(), nnz * 4, &has_value));
  scratch.resize((size_t)nrows * 4 + 16);
  WH_TRY(array(scratch.data(), (size_t)nrows * 4, &has));      // weight: decoded for validation, unused by this path
  if (nrows == 0) return WH_OK;
  return rows_append_host(ctx, out, (size_t)nrows, offset.data(), label.data(), index.data(), has_value ? value.data() : nullptr);
}

}  // namespace wh

using namespace wh;

extern "C" int wh_parse_crb(whContext* ctx, const char* data, size_t len, whRows** out) {
  WH_REQUIRE(ctx && out && (data || len == 0), "wh_parse_crb: null argument");
  DeviceGuard g(ctx->device);
  whRows* r = *out;
  const bool fresh = r == nullptr;
  if (fresh) WH_TRY(wh_rows_create(ctx, &r)); else WH_TRY(wh_rows_clear(r));
  struct Guard { whRows* r; ~Guard() { if (r) wh_rows_destroy(r); } } guard{fresh ? r : nullptr};
  const unsigned char* p = reinterpret_cast<const unsigned char*>(data);
  std::vector<unsigned char> rec, scratch;
  size_t pos = 0;
  while (pos + 8 <= len) {
    uint32_t magic, lrec;
    memcpy(&magic, p + pos, 4); memcpy(&lrec, p + pos + 4, 4);
    WH_REQUIRE(magic == kRecordMagic, "wh_parse_crb: no RecordIO magic at byte %zu (the chunk must start at a record)", pos);
    const uint32_t cflag = lrec >> 29, n = lrec & ((1u << 29) - 1);
    const size_t padded = ((size_t)n + 3) & ~(size_t)3;
    WH_REQUIRE(pos + 8 + n <= len, "wh_parse_crb: a record runs past the chunk (byte %zu)", pos);
    if (cflag == 0) {
      WH_TRY(crb_record(ctx, p + pos + 8, n, r, scratch));
    } else {
      // a payload that contained the magic: parts 1, 2.., 3, the magic re-inserted between them
      if (cflag == 1) rec.clear();
      else { const uint32_t m = kRecordMagic; rec.insert(rec.end(), reinterpret_cast<const unsigned char*>(&m), reinterpret_cast<const unsigned char*>(&m) + 4); }
      rec.insert(rec.end(), p + pos + 8, p + pos + 8 + n);
      if (cflag == 3) WH_TRY(crb_record(ctx, rec.data(), rec.size(), r, scratch));
    }
    pos += 8 + padded;
  }
  guard.r = nullptr;
  *out = r;
  return WH_OK;
}

extern "C" long wh_lz4_block_decode(const void* src, size_t n, void* dst, size_t cap) {
  return wh::lz4_block_decode(static_cast<const unsigned char*>(src), n, static_cast<unsigned char*>(dst), cap);
}
