// oracle/_ref harness, DiFacto part: a C-ABI around the reference's OWN headers,
// included unchanged from /root/reference/learn (nothing is copied into this repo).
// TEST INFRASTRUCTURE ONLY (see ref_linear.cc).
//
// Exercised reference code:
//   learn/difacto/loss.h:24-264        Loss::{Loss,Evaluate,CalcGrad}, Data::{Load,Save}
//   learn/difacto/async_sgd.h:135-425  AdaGradEntry, AdaGradHandle::{Push,Pull,Resize,UpdateW,UpdateV},
//                                      AsyncServer, AsyncWorker::ProcessMinibatch
#include "ref_common.h"
#include "ps.h"
#include "progress.h"
#include "async_sgd.h"

REF_HIDDEN void* operator new(size_t n) { void* p = calloc(1, n ? n : 1); if (!p) throw std::bad_alloc(); return p; }
REF_HIDDEN void* operator new[](size_t n) { void* p = calloc(1, n ? n : 1); if (!p) throw std::bad_alloc(); return p; }
REF_HIDDEN void operator delete(void* p) noexcept { free(p); }
REF_HIDDEN void operator delete[](void* p) noexcept { free(p); }
REF_HIDDEN void operator delete(void* p, size_t) noexcept { free(p); }
REF_HIDDEN void operator delete[](void* p, size_t) noexcept { free(p); }

namespace ps { uint64_t FLAGS_max_key = ~0ULL; }
namespace dmlc {
namespace difacto { int64_t ISGDHandle::new_w = 0; int64_t ISGDHandle::new_V = 0; }
Stream* Stream::Create(const char*, const char* const, bool) { return NULL; }
}  // namespace dmlc

using namespace dmlc;
using dmlc::difacto::Config;

namespace {
typedef ps::KVStoreSparse<float, difacto::AdaGradEntry, difacto::AdaGradHandle> Store;

void FillConf(Config* c, float l1, float l2, float eta, float beta, int l1_shrk, int dim,
              int threshold, float v_l2, int v_has_eta, float v_eta, int v_has_beta, float v_beta,
              float init_scale, float dropout, float clip, float norm, int nt) {
  c->set_lambda_l1(l1); c->set_lambda_l2(l2); c->set_lr_eta(eta); c->set_lr_beta(beta);
  c->set_l1_shrk(l1_shrk != 0); c->set_num_threads(nt); c->set_max_concurrency(1);
  if (dim > 0) {
    Config::Embedding* e = c->add_embedding();
    e->set_dim(dim); e->set_threshold(threshold); e->set_lambda_l2(v_l2);
    if (v_has_eta) e->set_lr_eta(v_eta);
    if (v_has_beta) e->set_lr_beta(v_beta);
    e->set_init_scale(init_scale); e->set_dropout(dropout); e->set_grad_clipping(clip);
    e->set_grad_normalization(norm);
  }
}

class Worker : public difacto::AsyncWorker {
 public:
  explicit Worker(const Config& c) : difacto::AsyncWorker(c) {}
  void Run(const RowBlock<uint64_t>& mb, bool train, int data_pass) {
    Workload wl; wl.type = train ? Workload::TRAIN : Workload::VAL; wl.data_pass = data_pass;
    wl.file.resize(1);
    ProcessMinibatch(mb, wl);
  }
};
struct Job { Config conf; difacto::AsyncServer* server; Worker* worker; ps::KVStore* store; };
}  // namespace

// Loss<float>(data, model, model_siz, conf); Evaluate; [CalcGrad]  (difacto/loss.h:24-158)
// prog = {objv, auc, objv_w, copc, count, new_ex, new_w, new_V}; grad (model_len) may be NULL.
REF_API void ref_difacto_loss(size_t rows, const size_t* offset, const unsigned* index,
                              const float* value, const float* label, const float* model,
                              size_t model_len, const int* model_siz, size_t nkeys, int dim, int nt,
                              float clip, float dropout, float norm, double* prog, float* grad) {
  Config c; FillConf(&c, 1, 0, .01f, 1, 1, dim, 0, 0, 0, 0, 0, 0, .01f, dropout, clip, norm, nt);
  RowBlock<unsigned> D; D.size = rows; D.offset = offset; D.index = index; D.value = value; D.label = label;
  std::vector<float> m(model, model + model_len);
  std::vector<int> ms(model_siz, model_siz + nkeys);
  difacto::Loss<float> loss(D, m, ms, c);
  difacto::Progress pg; loss.Evaluate(&pg);
  for (int i = 0; i < 8; ++i) prog[i] = pg.data[i];
  if (grad) { loss.CalcGrad(&m); memcpy(grad, m.data(), model_len * 4); }
}

REF_API void* ref_difacto_create(float l1, float l2, float eta, float beta, int l1_shrk, int dim,
                                 int threshold, float v_l2, int v_has_eta, float v_eta,
                                 int v_has_beta, float v_beta, float init_scale, float dropout,
                                 float clip, float norm, int nt, uint64_t max_key, unsigned seed) {
  ps::FLAGS_max_key = max_key;
  difacto::ISGDHandle::new_w = 0; difacto::ISGDHandle::new_V = 0;
  ps::FakeMonitor::Get().Clear();
  srand(seed);  // V init and gradient dropout draw from glibc rand() (async_sgd.h:254, loss.h:152)
  Job* j = new Job();
  FillConf(&j->conf, l1, l2, eta, beta, l1_shrk, dim, threshold, v_l2, v_has_eta, v_eta, v_has_beta,
           v_beta, init_scale, dropout, clip, norm, nt);
  j->server = new difacto::AsyncServer(j->conf);
  j->store = ps::KVStore::Current();
  j->worker = new Worker(j->conf);
  return j;
}
// One minibatch through AsyncWorker::ProcessMinibatch (difacto/async_sgd.h:363-425).
REF_API void ref_difacto_minibatch(void* h, const uint64_t* index, const size_t* offset,
                                   const float* value, const float* label, size_t rows, int train,
                                   int data_pass, double* prog) {
  Job* j = (Job*)h;
  ps::KVStore::Current() = j->store;
  ps::FakeMonitor& m = ps::FakeMonitor::Get();
  m.log.clear();
  RowBlock<uint64_t> b; b.size = rows; b.offset = offset; b.index = index; b.value = value; b.label = label;
  j->worker->Run(b, train != 0, data_pass);
  for (int i = 0; i < 8; ++i) prog[i] = 0;
  // worker vectors carry new_ex > 0; server vectors carry only new_w / new_V (async_sgd.h:68-75)
  for (size_t k = 0; k < m.log.size(); ++k) {
    if (m.log[k][5] > 0) { for (int i = 0; i < 6; ++i) prog[i] = m.log[k][i]; }
    prog[6] += m.log[k][6]; prog[7] += m.log[k][7];
  }
}
REF_API size_t ref_difacto_model_size(void* h) { return ((Store*)((Job*)h)->store)->data().size(); }
// Per key (ascending): fea_cnt, size; w/cg rows have stride (dim+1) / (dim+2), zero padded:
//   w  = {w0, V[0..dim)}        cg = {sqc_grad_0, z_0, sqc_grad_V[0..dim)}   (async_sgd.h:144-168)
REF_API void ref_difacto_model_get(void* h, int dim, uint64_t* keys, uint32_t* fea_cnt, int* size,
                                   float* w, float* cg) {
  size_t i = 0;
  for (auto& kv : ((Store*)((Job*)h)->store)->data()) {
    difacto::AdaGradEntry& e = kv.second;
    keys[i] = kv.first; fea_cnt[i] = e.fea_cnt; size[i] = e.size;
    float* wi = w + i * (dim + 1); float* ci = cg + i * (dim + 2);
    for (int k = 0; k < dim + 1; ++k) wi[k] = 0;
    for (int k = 0; k < dim + 2; ++k) ci[k] = 0;
    wi[0] = e.w_0(); ci[0] = e.sqc_grad_0(); ci[1] = e.z_0();
    for (int k = 1; k < e.size; ++k) { wi[k] = e.w[k]; ci[k + 1] = e.sqc_grad[k + 1]; }
    ++i;
  }
}
REF_API void ref_difacto_destroy(void* h) {
  Job* j = (Job*)h; delete j->worker; delete j->server; delete j->store; delete j;
}

// learn/difacto/progress.h:14-30 through the reference's own struct (8 values; carries ttl_ex, nnz_w, nnz_V).
REF_API int ref_difacto_progress_row(const double* data8, double* carry3, char* out, int cap) {
  dmlc::difacto::Progress p;
  p.data.assign(data8, data8 + 8);
  p.ttl_ex = carry3[0]; p.nnz_w = carry3[1]; p.nnz_V = carry3[2];
  std::string s = p.PrintStr();
  carry3[0] = p.ttl_ex; carry3[1] = p.nnz_w; carry3[2] = p.nnz_V;
  snprintf(out, cap, "%s", s.c_str());
  return (int)s.size();
}
REF_API int ref_difacto_progress_head(char* out, int cap) {
  return snprintf(out, cap, "%s", dmlc::difacto::Progress::HeadStr().c_str());
}
