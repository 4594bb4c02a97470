// dmlc::Localizer<I>, same interface as learn/base/localizer.h:28-95, forwarding to wh_localize.
#pragma once
#include <limits>
#include <type_traits>
#include <vector>

#include "data/row_block.h"
#include "dmlc/data.h"
#include "ps.h"
#include "wh_runtime.h"

namespace dmlc {

/// learn/base/localizer.h:16-26 (a nibble reversal)
inline uint64_t ReverseBytes(uint64_t x) { return wh_reverse_bytes(x); }

template <typename I>
class Localizer {
  static_assert(sizeof(I) == 8, "the GPU localizer handles 64-bit feature ids (ps::Key)");
 public:
  explicit Localizer(int nthreads = 2) { (void)nthreads; }
  ~Localizer() { Clear(); }

  /// learn/base/localizer.h:42-51
  template <typename C = unsigned>
  void Localize(const RowBlock<I>& blk, data::RowBlockContainer<unsigned>* localized,
                std::vector<I>* uniq_idx = NULL, std::vector<C>* idx_frq = NULL) {
    std::vector<I> tmp;
    CountUniqIndex<C>(blk, uniq_idx ? uniq_idx : &tmp, idx_frq);
    RemapIndex(blk, uniq_idx ? *uniq_idx : tmp, localized);
    Clear();
  }

  /// learn/base/localizer.h:96-167
  template <typename C>
  void CountUniqIndex(const RowBlock<I>& blk, std::vector<I>* uniq_idx, std::vector<C>* idx_frq) {
    if (blk.size == 0) return;
    Clear();
    whRowBlock rb = View(blk);
    WH_CHECK(wh_localize(wh::Ctx(), &rb, WH_HOST, ps::FLAGS_max_key, WH_LOC_AUTO | (idx_frq ? WH_LOC_WITH_COUNTS : 0), &batch_));
    const size_t U = wh_batch_num_uniq(batch_);
    uniq_idx->resize(U);
    std::vector<float> cf;
    std::vector<uint32_t> cu;
    if (idx_frq) {
      if (std::is_integral<C>::value) cu.resize(U); else cf.resize(U);
    }
    WH_CHECK(wh_batch_copy(wh::Ctx(), batch_, reinterpret_cast<uint64_t*>(uniq_idx->data()),
                           cu.empty() ? NULL : cu.data(), cf.empty() ? NULL : cf.data(), NULL, NULL, NULL, NULL));
    if (idx_frq) {
      idx_frq->resize(U);
      const unsigned cmax = std::is_integral<C>::value ? static_cast<unsigned>(std::numeric_limits<C>::max()) : 0;
      for (size_t i = 0; i < U; ++i)
        (*idx_frq)[i] = std::is_integral<C>::value ? static_cast<C>(cu[i] < cmax ? cu[i] : cmax) : static_cast<C>(cf[i]);
    }
    own_dict_ = uniq_idx;
  }

  /// learn/base/localizer.h:169-221
  void RemapIndex(const RowBlock<I>& blk, const std::vector<I>& idx_dict, data::RowBlockContainer<unsigned>* o) {
    if (blk.size == 0 || idx_dict.empty()) return;
    whBatch* b = batch_;
    bool mine = false;
    if (!(b && own_dict_ == &idx_dict)) {   // a foreign dictionary: unmatched entries are dropped
      whRowBlock rb = View(blk);
      WH_CHECK(wh_localize_with_dict(wh::Ctx(), &rb, WH_HOST, ps::FLAGS_max_key,
                                     reinterpret_cast<const uint64_t*>(idx_dict.data()), idx_dict.size(), &b));
      mine = true;
    }
    const size_t n = wh_batch_nnz(b);
    o->offset.resize(blk.size + 1);
    o->index.resize(n);
    if (blk.value) o->value.resize(n);
    if (blk.label) o->label.resize(blk.size);
    WH_CHECK(wh_batch_copy(wh::Ctx(), b, NULL, NULL, NULL, o->offset.data(), o->index.data(),
                           (blk.value && n) ? o->value.data() : NULL, blk.label ? o->label.data() : NULL));
    o->max_index = static_cast<unsigned>(idx_dict.size() - 1);
    if (mine) wh_batch_destroy(b);
  }

  void Clear() {
    if (batch_) wh_batch_destroy(batch_);
    batch_ = nullptr; own_dict_ = nullptr;
  }

 private:
  static whRowBlock View(const RowBlock<I>& blk) {
    whRowBlock rb;
    rb.size = blk.size; rb.offset = blk.offset; rb.label = blk.label; rb.weight = blk.weight;
    rb.index = reinterpret_cast<const uint64_t*>(blk.index); rb.value = blk.value;
    return rb;
  }
  whBatch* batch_ = nullptr;       // the "temporal results" CountUniqIndex keeps for RemapIndex
  const void* own_dict_ = nullptr;
};

}  // namespace dmlc
