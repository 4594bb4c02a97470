// dmlc::data::MinibatchIter<uint64_t> with the reference's constructor and methods
// (learn/base/minibatch_iter.h:29-148), reading text with the GPU reader: chunks of whole lines go to the device
// as they are (wh_parse_text), and Next() assembles exactly `minibatch_size` rows across chunk borders in a device
// container (wh_rows_*). The host side only cuts the file at line ends and runs the reference's rand()-driven
// bookkeeping -- std::random_shuffle's draw order (libstdc++) and the negative-sampling test -- on 4 bytes per row.
//
//   Value()        the minibatch as a host RowBlock, as the reference returns it (copied back on demand)
//   DeviceValue()  the same minibatch as DEVICE pointers + its entry count, for
//                  AsgdWorker::ProcessMinibatch(const DeviceMinibatch&) / wh_worker_submit(..., WH_DEVICE, ...)
//
// Formats: libsvm, criteo, criteo_test, adfea (crb needs lz4 + recordio: LOG(FATAL) equivalent).
#pragma once
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "dmlc/data.h"
#include "wh_runtime.h"

namespace dmlc {
namespace data {

struct DeviceMinibatch {
  whRowBlock blk;   // device pointers
  size_t nnz = 0;
};

namespace detail {
inline bool IsBreak(int c) { return c == '\n' || c == '\r'; }

// One part of one text file, served in chunks of whole lines (InputSplit "text": a part owns the lines that start
// inside its byte range; both ends move forward to the next line start).
class TextSplit {
 public:
  TextSplit(const char* uri, unsigned part, unsigned nparts, size_t chunk_bytes) : chunk_(chunk_bytes) {
    f_ = fopen(uri, "rb");
    if (!f_) { fprintf(stderr, "[FATAL] cannot open %s\n", uri); abort(); }
    fseek(f_, 0, SEEK_END);
    total_ = (size_t)ftell(f_);
    begin_ = Align(total_ * part / nparts);
    end_ = Align(total_ * (part + 1) / nparts);
    BeforeFirst();
  }
  ~TextSplit() { if (f_) fclose(f_); }
  void BeforeFirst() { pos_ = begin_; bytes_read_ = 0; }
  size_t BytesRead() const { return bytes_read_; }
  // false at the end of the part; otherwise buf holds the next chunk
  bool NextChunk(std::vector<char>* buf) {
    if (pos_ >= end_) return false;
    size_t want = chunk_ < end_ - pos_ ? chunk_ : end_ - pos_;
    buf->resize(want);
    fseek(f_, (long)pos_, SEEK_SET);
    if (fread(buf->data(), 1, want, f_) != want) { fprintf(stderr, "[FATAL] short read\n"); abort(); }
    if (pos_ + want < end_) {                     // finish the current line and take its break(s)
      bool in_break = IsBreak(buf->back());
      int c;
      while ((c = fgetc(f_)) != EOF) {
        if (IsBreak(c)) in_break = true;
        else if (in_break) break;
        buf->push_back((char)c);
      }
    }
    pos_ += buf->size();
    bytes_read_ += buf->size();
    return true;
  }
 private:
  size_t Align(size_t x) {
    if (x == 0 || x >= total_) return x < total_ ? x : total_;
    fseek(f_, (long)x, SEEK_SET);
    int c;
    while ((c = fgetc(f_)) != EOF && !IsBreak(c)) ++x;      // rest of the current line
    if (c == EOF) return total_;
    ++x;
    while ((c = fgetc(f_)) != EOF && IsBreak(c)) ++x;       // its break(s)
    return x;
  }
  FILE* f_ = nullptr;
  size_t total_ = 0, begin_ = 0, end_ = 0, pos_ = 0, chunk_, bytes_read_ = 0;
};
}  // namespace detail

template <typename IndexType>
class MinibatchIter {
  static_assert(sizeof(IndexType) == 8, "the GPU reader produces 64-bit feature ids");
 public:
  MinibatchIter(const char* uri, unsigned part_index, unsigned num_parts, const char* type, unsigned minibatch_size,
                unsigned shuf_buf = 0, float negative_sampling = 1.0, size_t chunk_bytes = (size_t)8 << 20, int ring = 4)
      : mb_size_(minibatch_size), shuf_buf_(shuf_buf), negative_sampling_(negative_sampling) {
    if (!strcmp(type, "libsvm")) fmt_ = WH_FMT_LIBSVM;
    else if (!strcmp(type, "criteo")) fmt_ = WH_FMT_CRITEO;
    else if (!strcmp(type, "criteo_test")) fmt_ = WH_FMT_CRITEO_TEST;
    else if (!strcmp(type, "adfea")) fmt_ = WH_FMT_ADFEA;
    else { fprintf(stderr, "[FATAL] unknown datatype %s (libsvm, criteo, criteo_test, adfea are built)\n", type); abort(); }
    if (shuf_buf) {
      if (!(shuf_buf > minibatch_size)) { fprintf(stderr, "[FATAL] Check failed: shuf_buf > minibatch_size\n"); abort(); }
      buf_reader_ = new MinibatchIter(uri, part_index, num_parts, type, shuf_buf, 0, 1.0f, chunk_bytes, 2);
    } else {
      split_ = new detail::TextSplit(uri, part_index, num_parts, chunk_bytes);
    }
    // a submitted minibatch is read asynchronously by the worker: hand out `ring` containers in turn
    ring_.resize(ring < 1 ? 1 : ring, nullptr);
    for (auto& r : ring_) WH_CHECK(wh_rows_create(wh::Ctx(), &r));
  }
  virtual ~MinibatchIter() {
    for (auto r : ring_) wh_rows_destroy(r);
    if (chunk_rows_) wh_rows_destroy(chunk_rows_);
    delete split_;
    delete buf_reader_;
  }
  void BeforeFirst() {
    if (split_) split_->BeforeFirst();
    if (buf_reader_) buf_reader_->BeforeFirst();
    cur_ = nullptr; start_ = end_ = 0;
  }
  bool Next() {
    whContext* ctx = wh::Ctx();
    out_ = ring_[turn_];
    turn_ = (turn_ + 1) % ring_.size();
    WH_CHECK(wh_rows_clear(out_));
    host_valid_ = false;
    size_t have = 0;
    while (have < mb_size_) {
      if (start_ == end_) {
        if (shuf_buf_ == 0) {
          // no random shuffle: the next non-empty parsed chunk
          cur_ = nullptr;
          while (split_->NextChunk(&text_)) {
            WH_CHECK(wh_parse_text(ctx, fmt_, text_.data(), text_.size(), WH_HOST, &chunk_rows_));
            if (wh_rows_size(chunk_rows_)) { cur_ = chunk_rows_; break; }
          }
          if (!cur_) { start_ = end_ = 0; break; }
        } else {
          if (!buf_reader_->Next()) { cur_ = nullptr; start_ = end_ = 0; break; }
          cur_ = buf_reader_->out_;
          const size_t n = wh_rows_size(cur_);
          if (rdp_.size() != n) { rdp_.resize(n); for (size_t i = 0; i < n; ++i) rdp_[i] = (unsigned)i; }
          for (size_t i = 1; i < n; ++i) {          // std::random_shuffle as libstdc++ draws it (:92)
            const size_t j = (size_t)rand() % (i + 1);
            if (i != j) std::swap(rdp_[i], rdp_[j]);
          }
          if (negative_sampling_ < 1.0) {
            cur_label_.resize(n);
            WH_CHECK(wh_rows_copy(ctx, cur_, nullptr, cur_label_.data(), nullptr, nullptr));
          }
        }
        start_ = 0; end_ = wh_rows_size(cur_);
      }
      const size_t len = std::min(end_ - start_, (size_t)mb_size_ - have);      // :98
      if (shuf_buf_ == 0) {
        WH_CHECK(wh_rows_append(ctx, out_, cur_, start_, len, nullptr));
        have += len;
      } else {
        keep_.clear();
        for (size_t i = start_; i < start_ + len; ++i) {                          // :102-110
          const unsigned j = rdp_[i];
          if (negative_sampling_ < 1.0 && cur_label_[j] <= 0 &&
              (float)rand() / (float)RAND_MAX > 1 - negative_sampling_) continue;
          keep_.push_back(j);
        }
        if (!keep_.empty()) WH_CHECK(wh_rows_append(ctx, out_, cur_, 0, keep_.size(), keep_.data()));
        have += keep_.size();
      }
      start_ += len;
    }
    WH_CHECK(wh_rows_drop_unit_values(ctx, out_));                                // :116-118
    WH_CHECK(wh_sync(ctx));
    return wh_rows_size(out_) > 0;
  }
  size_t BytesRead() const { return split_ ? split_->BytesRead() : buf_reader_->BytesRead(); }

  /// the minibatch on the device (valid until `ring` further calls of Next)
  DeviceMinibatch DeviceValue() const {
    DeviceMinibatch d;
    WH_CHECK(wh_rows_view(out_, &d.blk));
    d.nnz = wh_rows_nnz(out_);
    return d;
  }
  /// the minibatch as the reference returns it: host arrays (copied back on first use)
  const RowBlock<IndexType>& Value() {
    if (!host_valid_) {
      const size_t rows = wh_rows_size(out_), nnz = wh_rows_nnz(out_);
      h_offset_.resize(rows + 1); h_label_.resize(rows); h_index_.resize(nnz);
      h_value_.resize(wh_rows_has_value(out_) ? nnz : 0);
      WH_CHECK(wh_rows_copy(wh::Ctx(), out_, h_offset_.data(), h_label_.data(), h_index_.data(),
                            h_value_.empty() ? nullptr : h_value_.data()));
      host_blk_.size = rows;
      host_blk_.offset = h_offset_.data();
      host_blk_.label = h_label_.data();
      host_blk_.weight = nullptr;
      host_blk_.index = reinterpret_cast<const IndexType*>(h_index_.data());
      host_blk_.value = h_value_.empty() ? nullptr : h_value_.data();
      host_valid_ = true;
    }
    return host_blk_;
  }

 private:
  unsigned mb_size_, shuf_buf_;
  float negative_sampling_;
  int fmt_ = 0;
  detail::TextSplit* split_ = nullptr;
  MinibatchIter* buf_reader_ = nullptr;
  std::vector<whRows*> ring_;
  size_t turn_ = 0;
  whRows* out_ = nullptr;
  whRows* chunk_rows_ = nullptr;   // the parsed chunk (buffers reused chunk after chunk)
  whRows* cur_ = nullptr;          // where rows are taken from: chunk_rows_ or the buffer reader's minibatch
  size_t start_ = 0, end_ = 0;
  std::vector<char> text_;
  std::vector<unsigned> rdp_;
  std::vector<uint32_t> keep_;
  std::vector<float> cur_label_;
  bool host_valid_ = false;
  RowBlock<IndexType> host_blk_;
  std::vector<size_t> h_offset_;
  std::vector<float> h_label_, h_value_;
  std::vector<uint64_t> h_index_;
};

}  // namespace data
}  // namespace dmlc
