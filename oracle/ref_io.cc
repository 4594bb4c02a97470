// oracle/_ref/libwh_ref_io.so: the reference's own reader -- learn/base/minibatch_iter.h with
// learn/base/criteo_parser.h and adfea_parser.h, compiled unchanged from /root/reference -- behind the
// stand-ins in shims_io/ (InputSplit, ParserImpl, libsvm parser, CityHash64) and shims/. TEST INFRASTRUCTURE ONLY.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <city.h>
#include "base/minibatch_iter.h"
#include "ref_common.h"

typedef dmlc::data::MinibatchIter<uint64_t> Iter;

REF_API uint64_t ref_cityhash64(const char* s, size_t len) { return CityHash64(s, len); }
REF_API void* ref_iter_open(const char* uri, unsigned part, unsigned nparts, const char* type, unsigned mb_size,
                            unsigned shuf_buf, float neg_sampling) {
  return new Iter(uri, part, nparts, type, mb_size, shuf_buf, neg_sampling);
}
REF_API void ref_iter_close(void* it) { delete static_cast<Iter*>(it); }
REF_API void ref_iter_before_first(void* it) { static_cast<Iter*>(it)->BeforeFirst(); }
// returns 1 and the sizes of the minibatch, or 0 at the end of the data
REF_API int ref_iter_next(void* it, size_t* rows, size_t* nnz, int* has_value) {
  Iter* I = static_cast<Iter*>(it);
  if (!I->Next()) return 0;
  const dmlc::RowBlock<uint64_t>& b = I->Value();
  *rows = b.size; *nnz = b.offset[b.size] - b.offset[0]; *has_value = b.value != NULL;
  return 1;
}
REF_API void ref_iter_copy(void* it, uint64_t* offset, float* label, uint64_t* index, float* value) {
  const dmlc::RowBlock<uint64_t>& b = static_cast<Iter*>(it)->Value();
  size_t n = b.offset[b.size] - b.offset[0];
  for (size_t i = 0; i <= b.size; ++i) offset[i] = b.offset[i] - b.offset[0];
  memcpy(label, b.label, b.size * sizeof(float));
  memcpy(index, b.index + b.offset[0], n * sizeof(uint64_t));
  if (b.value && value) memcpy(value, b.value + b.offset[0], n * sizeof(float));
}
REF_API size_t ref_iter_bytes_read(void* it) { return static_cast<Iter*>(it)->BytesRead(); }
