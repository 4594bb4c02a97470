// oracle/_ref harness, linear part: a C-ABI around the reference's OWN headers,
// included unchanged from /root/reference/learn (nothing is copied into this repo).
// TEST INFRASTRUCTURE ONLY: only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load the resulting oracle/_ref/libwh_ref.so.
//
// Exercised reference code:
//   learn/base/localizer.h:16-221      ReverseBytes, Localizer::{Localize,CountUniqIndex,RemapIndex}
//   learn/base/parallel_sort.h:15-41   ParallelSort
//   learn/base/spmv.h:72-119           SpMV::Times / TransTimes
//   learn/base/spmm.h:23-123           SpMM::Times / TransTimes
//   learn/base/binary_class_evaluation.h:17-74  AUC / Accuracy / LogLoss / LogitObjv
//   learn/linear/loss.h:20-182         ScalarLoss, LogitLoss, SquareHingeLoss, CreateLoss
//   learn/linear/penalty.h:36-41       L1L2::Solve
//   learn/linear/async_sgd.h:21-305    handles, AsgdServer, AsgdWorker::ProcessMinibatch
#include "ref_common.h"
#include "ps.h"
#include "async_sgd.h"
#include "base/spmm.h"

REF_HIDDEN void* operator new(size_t n) { void* p = calloc(1, n ? n : 1); if (!p) throw std::bad_alloc(); return p; }
REF_HIDDEN void* operator new[](size_t n) { void* p = calloc(1, n ? n : 1); if (!p) throw std::bad_alloc(); return p; }
REF_HIDDEN void operator delete(void* p) noexcept { free(p); }
REF_HIDDEN void operator delete[](void* p) noexcept { free(p); }
REF_HIDDEN void operator delete(void* p, size_t) noexcept { free(p); }
REF_HIDDEN void operator delete[](void* p, size_t) noexcept { free(p); }

namespace ps { uint64_t FLAGS_max_key = ~0ULL; }
namespace dmlc {
namespace linear { int64_t ISGDHandle::new_w = 0; }
Stream* Stream::Create(const char*, const char* const, bool) { return NULL; }
}  // namespace dmlc

using namespace dmlc;
using dmlc::linear::Config;

namespace {
RowBlock<uint64_t> MakeBlk(const uint64_t* index, const size_t* offset, const float* value,
                           const float* label, size_t rows) {
  RowBlock<uint64_t> b;
  b.size = rows; b.offset = offset; b.index = index; b.value = value; b.label = label;
  return b;
}
RowBlock<unsigned> MakeLocal(size_t rows, const size_t* offset, const unsigned* index,
                             const float* value, const float* label) {
  RowBlock<unsigned> b;
  b.size = rows; b.offset = offset; b.index = index; b.value = value; b.label = label;
  return b;
}
struct LocResult {
  data::RowBlockContainer<unsigned> data;
  std::vector<uint64_t> uniq;
  std::vector<unsigned> cnt_u;
  std::vector<float> cnt_f;
};
}  // namespace

REF_API uint64_t ref_reverse_bytes(uint64_t x) { return ReverseBytes(x); }

// Localizer<uint64>::Localize. count_mode: 0 none, 1 unsigned, 2 float (DiFacto, difacto/async_sgd.h:366)
REF_API void* ref_localize(const uint64_t* index, const size_t* offset, const float* value,
                           const float* label, size_t rows, uint64_t max_key, int nt,
                           int count_mode) {
  ps::FLAGS_max_key = max_key;
  LocResult* r = new LocResult();
  Localizer<uint64_t> lc(nt);
  RowBlock<uint64_t> blk = MakeBlk(index, offset, value, label, rows);
  if (count_mode == 2) lc.Localize<float>(blk, &r->data, &r->uniq, &r->cnt_f);
  else if (count_mode == 1) lc.Localize<unsigned>(blk, &r->data, &r->uniq, &r->cnt_u);
  else lc.Localize<unsigned>(blk, &r->data, &r->uniq);
  return r;
}
// CountUniqIndex on blk, then RemapIndex against an external ascending dictionary
// (unmatched entries are dropped: localizer.h:179-213).
REF_API void* ref_remap(const uint64_t* index, const size_t* offset, const float* value,
                        const float* label, size_t rows, uint64_t max_key, int nt,
                        const uint64_t* dict, size_t ndict) {
  ps::FLAGS_max_key = max_key;
  LocResult* r = new LocResult();
  Localizer<uint64_t> lc(nt);
  RowBlock<uint64_t> blk = MakeBlk(index, offset, value, label, rows);
  lc.CountUniqIndex<unsigned>(blk, &r->uniq, &r->cnt_u);
  std::vector<uint64_t> d(dict, dict + ndict);
  lc.RemapIndex(blk, d, &r->data);
  return r;
}
REF_API size_t ref_loc_nuniq(void* h) { return ((LocResult*)h)->uniq.size(); }
REF_API size_t ref_loc_nnz(void* h) { return ((LocResult*)h)->data.index.size(); }
REF_API size_t ref_loc_rows(void* h) { return ((LocResult*)h)->data.offset.size() - 1; }
REF_API int ref_loc_has_value(void* h) { return !((LocResult*)h)->data.value.empty(); }
REF_API int ref_loc_has_label(void* h) { return !((LocResult*)h)->data.label.empty(); }
REF_API uint32_t ref_loc_max_index(void* h) { return ((LocResult*)h)->data.max_index; }
REF_API void ref_loc_get(void* h, uint64_t* uniq, uint32_t* cnt_u, float* cnt_f, size_t* offset,
                         uint32_t* index, float* value, float* label) {
  LocResult* r = (LocResult*)h;
  if (uniq) memcpy(uniq, r->uniq.data(), r->uniq.size() * 8);
  if (cnt_u) memcpy(cnt_u, r->cnt_u.data(), r->cnt_u.size() * 4);
  if (cnt_f) memcpy(cnt_f, r->cnt_f.data(), r->cnt_f.size() * 4);
  if (offset) memcpy(offset, r->data.offset.data(), r->data.offset.size() * sizeof(size_t));
  if (index) memcpy(index, r->data.index.data(), r->data.index.size() * 4);
  if (value) memcpy(value, r->data.value.data(), r->data.value.size() * 4);
  if (label) memcpy(label, r->data.label.data(), r->data.label.size() * 4);
}
REF_API void ref_loc_free(void* h) { delete (LocResult*)h; }

REF_API void ref_spmv_times(size_t rows, const size_t* offset, const unsigned* index,
                            const float* value, const float* x, float* y, int nt) {
  SpMV::Times<float>(MakeLocal(rows, offset, index, value, NULL), x, y, nt);
}
REF_API void ref_spmv_trans_times(size_t rows, const size_t* offset, const unsigned* index,
                                  const float* value, const float* x, float* y, size_t ysize,
                                  int nt) {
  SpMV::TransTimes<float>(MakeLocal(rows, offset, index, value, NULL), x, y, ysize, nt);
}
REF_API void ref_spmm_times(size_t rows, const size_t* offset, const unsigned* index,
                            const float* value, const float* x, size_t xsize, float* y,
                            size_t ysize, int nt) {
  std::vector<float> xv(x, x + xsize), yv(y, y + ysize);
  SpMM::Times<float>(MakeLocal(rows, offset, index, value, NULL), xv, &yv, nt);
  memcpy(y, yv.data(), ysize * 4);
}
REF_API void ref_spmm_trans_times(size_t rows, const size_t* offset, const unsigned* index,
                                  const float* value, const float* x, size_t xsize, float p,
                                  const float* z, float* y, size_t ysize, int nt) {
  std::vector<float> xv(x, x + xsize), yv(y, y + ysize), zv;
  if (z) zv.assign(z, z + ysize);
  SpMM::TransTimes<float>(MakeLocal(rows, offset, index, value, NULL), xv, p, zv, &yv, nt);
  memcpy(y, yv.data(), ysize * 4);
}

// out = {AUC, Accuracy(0), LogLoss, LogitObjv, Copc}
REF_API void ref_bin_eval(const float* label, const float* predict, size_t n, int nt, float* out) {
  BinClassEval<float> ev(label, predict, n, nt);
  out[0] = ev.AUC(); out[1] = ev.Accuracy(0); out[2] = ev.LogLoss(); out[3] = ev.LogitObjv();
  out[4] = ev.Copc();
}

REF_API float ref_l1l2_solve(float l1, float l2, float z, float eta) {
  linear::L1L2<float> p; p.set_lambda1(l1); p.set_lambda2(l2); return p.Solve(z, eta);
}

// CreateLoss(loss)->Init(data, w, nt); Evaluate(prog); [CalcGrad(grad)]  (loss.h:32-49,172-182)
// prog = {objv, acc, auc, count, new_ex, new_w}; grad (size U) may be NULL to skip CalcGrad.
REF_API void ref_linear_loss(int loss_type, size_t rows, const size_t* offset, const unsigned* index,
                             const float* value, const float* label, const float* w, size_t U,
                             int nt, double* prog, float* grad) {
  RowBlock<unsigned> D = MakeLocal(rows, offset, index, value, label);
  std::vector<float> wv(w, w + U);
  linear::ScalarLoss<float>* loss = linear::CreateLoss<float>((Config::Loss)loss_type);
  loss->Init(D, wv, nt);
  linear::Progress pg; loss->Evaluate(&pg);
  for (int i = 0; i < 6; ++i) prog[i] = pg.data[i];
  if (grad) { loss->CalcGrad(&wv); memcpy(grad, wv.data(), U * 4); }
  delete loss;
}

// One push message applied by the real handle to caller-held entry arrays.
// algo: Config::Algo. t_inout: SGDHandle::t (starts at 1, async_sgd.h:100). Returns delta |w|_0.
REF_API int64_t ref_handle_push(int algo, float l1, float l2, int has_eta, float eta, int has_beta,
                                float beta, size_t n, const float* grad, float* w, float* z,
                                float* cg, int* t_inout) {
  linear::ISGDHandle::new_w = 0;
  auto cfg = [&](linear::ISGDHandle& h) {
    h.penalty.set_lambda1(l1); h.penalty.set_lambda2(l2);
    if (has_eta) h.alpha = eta;
    if (has_beta) h.beta = beta;
  };
  if (algo == Config::SGD) {
    linear::SGDHandle h; cfg(h); h.t = *t_inout;
    h.Start(true, 0, 0, NULL);
    for (size_t i = 0; i < n; ++i) {
      linear::SGDEntry e; e.w = w[i];
      h.Push(0, ps::Blob<const float>(grad + i, 1), e);
      w[i] = e.w;
    }
    *t_inout = h.t;
  } else if (algo == Config::ADAGRAD) {
    linear::AdaGradHandle h; cfg(h);
    h.Start(true, 0, 0, NULL);
    for (size_t i = 0; i < n; ++i) {
      linear::AdaGradEntry e; e.w = w[i]; e.sq_cum_grad = cg[i];
      h.Push(0, ps::Blob<const float>(grad + i, 1), e);
      w[i] = e.w; cg[i] = e.sq_cum_grad;
    }
  } else {
    linear::FTRLHandle h; cfg(h);
    h.Start(true, 0, 0, NULL);
    for (size_t i = 0; i < n; ++i) {
      linear::FTRLEntry e; e.w = w[i]; e.z = z[i]; e.sq_cum_grad = cg[i];
      h.Push(0, ps::Blob<const float>(grad + i, 1), e);
      w[i] = e.w; z[i] = e.z; cg[i] = e.sq_cum_grad;
    }
  }
  return linear::ISGDHandle::new_w;
}

// ---- full path: AsgdServer + AsgdWorker::ProcessMinibatch over the in-process ps stand-in ----
namespace {
class Worker : public linear::AsgdWorker {
 public:
  explicit Worker(const Config& c) : linear::AsgdWorker(c) {}
  void Run(const RowBlock<uint64_t>& mb, bool train) {
    Workload wl; wl.type = train ? Workload::TRAIN : Workload::VAL; wl.data_pass = 0;
    wl.file.resize(1);
    ProcessMinibatch(mb, wl);
  }
};
struct LinearJob {
  Config conf;
  linear::AsgdServer* server;
  Worker* worker;
  ps::KVStore* store;
};
}  // namespace

REF_API void* ref_linear_create(int algo, int loss, float l1, float l2, int has_eta, float eta,
                                int has_beta, float beta, uint64_t max_key) {
  ps::FLAGS_max_key = max_key;
  linear::ISGDHandle::new_w = 0;
  ps::FakeMonitor::Get().Clear();
  LinearJob* j = new LinearJob();
  j->conf.set_algo((Config::Algo)algo);
  j->conf.set_loss((Config::Loss)loss);
  j->conf.set_lambda_l1(l1); j->conf.set_lambda_l2(l2);
  if (has_eta) j->conf.set_lr_eta(eta);
  if (has_beta) j->conf.set_lr_beta(beta);
  j->conf.set_max_concurrency(1);
  j->server = new linear::AsgdServer(j->conf);
  j->store = ps::KVStore::Current();
  j->worker = new Worker(j->conf);
  return j;
}
// Runs one minibatch through the reference's ProcessMinibatch (async_sgd.h:240-288).
// prog receives the worker's progress vector {objv, acc, auc, count, new_ex, new_w=0}.
REF_API void ref_linear_minibatch(void* h, const uint64_t* index, const size_t* offset,
                                  const float* value, const float* label, size_t rows, int train,
                                  double* prog) {
  LinearJob* j = (LinearJob*)h;
  ps::KVStore::Current() = j->store;
  ps::FakeMonitor& m = ps::FakeMonitor::Get();
  m.log.clear();
  j->worker->Run(MakeBlk(index, offset, value, label, rows), train != 0);
  // Vectors arrive from the server handle's Finish() after the pull and after the push
  // ({0,..,new_w}, async_sgd.h:26-33) and from the worker's callback (new_ex > 0, :262).
  for (int i = 0; i < 6; ++i) prog[i] = 0;
  for (size_t k = 0; k < m.log.size(); ++k) {
    if (m.log[k][4] > 0) { for (int i = 0; i < 5; ++i) prog[i] = m.log[k][i]; }
    prog[5] += m.log[k][5];
  }
}
REF_API size_t ref_linear_model_size(void* h) {
  LinearJob* j = (LinearJob*)h;
  int algo = j->conf.algo();
  if (algo == Config::SGD)
    return ((ps::KVStoreSparse<float, linear::SGDEntry, linear::SGDHandle>*)j->store)->data().size();
  if (algo == Config::ADAGRAD)
    return ((ps::KVStoreSparse<float, linear::AdaGradEntry, linear::AdaGradHandle>*)j->store)->data().size();
  return ((ps::KVStoreSparse<float, linear::FTRLEntry, linear::FTRLHandle>*)j->store)->data().size();
}
// keys ascending; z / cg are filled where the entry type has them (else left untouched).
REF_API void ref_linear_model_get(void* h, uint64_t* keys, float* w, float* z, float* cg) {
  LinearJob* j = (LinearJob*)h;
  int algo = j->conf.algo();
  size_t i = 0;
  if (algo == Config::SGD) {
    for (auto& kv : ((ps::KVStoreSparse<float, linear::SGDEntry, linear::SGDHandle>*)j->store)->data()) {
      keys[i] = kv.first; w[i] = kv.second.w; ++i;
    }
  } else if (algo == Config::ADAGRAD) {
    for (auto& kv : ((ps::KVStoreSparse<float, linear::AdaGradEntry, linear::AdaGradHandle>*)j->store)->data()) {
      keys[i] = kv.first; w[i] = kv.second.w; if (cg) cg[i] = kv.second.sq_cum_grad; ++i;
    }
  } else {
    for (auto& kv : ((ps::KVStoreSparse<float, linear::FTRLEntry, linear::FTRLHandle>*)j->store)->data()) {
      keys[i] = kv.first; w[i] = kv.second.w; if (z) z[i] = kv.second.z;
      if (cg) cg[i] = kv.second.sq_cum_grad; ++i;
    }
  }
}
REF_API void ref_linear_destroy(void* h) {
  LinearJob* j = (LinearJob*)h;
  delete j->worker; delete j->server; delete j->store; delete j;
}
REF_API int ref_num_procs(void) { return omp_get_num_procs(); }

// learn/linear/progress.h:14-27 through the reference's own struct: one table row for the merged progress
// `data6`, carrying ttl_ex / nnz_w across calls. Returns the length written to out (0 = "no progress").
REF_API int ref_linear_progress_row(const double* data6, double* ttl_ex, double* nnz_w, char* out, int cap) {
  dmlc::linear::Progress p;
  p.data.assign(data6, data6 + 6);
  p.ttl_ex = *ttl_ex; p.nnz_w = *nnz_w;
  std::string s = p.PrintStr();
  *ttl_ex = p.ttl_ex; *nnz_w = p.nnz_w;
  snprintf(out, cap, "%s", s.c_str());
  return (int)s.size();
}
REF_API int ref_linear_progress_head(char* out, int cap) {
  return snprintf(out, cap, "%s", dmlc::linear::Progress::HeadStr().c_str());
}
