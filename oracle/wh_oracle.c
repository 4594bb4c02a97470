/*
 * wh_oracle.c -- CPU restatement ("port") of wormhole's sparse minibatch
 * training hot path, in plain C.
 *
 * TEST INFRASTRUCTURE ONLY. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this. It is the checker, never
 * the thing shipped: the product path (wormhole_b200/csrc) has no CPU fallback.
 *
 * Parity status: PINNED. Every function below is checked bit-for-bit (float
 * outputs included, at nt<=2 where the reference's OpenMP reduction order is
 * deterministic) against oracle/_ref/ -- the reference's own headers compiled
 * unchanged -- by tests/test_oracle_vs_ref.py in the build container, and
 * against the golden vectors under tests/golden/ (generated from oracle/_ref by
 * tests/golden/make_golden.py) everywhere else. The one part the reference tree
 * does not contain -- the ps-lite v1 server loop that applies a handle to a
 * message -- is restated from its call sites (see oracle/shims/ps.h) and is
 * unpinned by any in-tree reference test.
 *
 * Each function cites the reference lines it follows (paths relative to
 * /root/reference/).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------ keys -- */

/* learn/base/localizer.h:16-26  ReverseBytes: despite the name it reverses the
 * sixteen 4-bit nibbles (32/16/8/4-bit swaps). */
ORC_API uint64_t orc_reverse_bytes(uint64_t x) {
  x = x << 32 | x >> 32;
  x = (x & 0x0000FFFF0000FFFFULL) << 16 | (x & 0xFFFF0000FFFF0000ULL) >> 16;
  x = (x & 0x00FF00FF00FF00FFULL) << 8 | (x & 0xFF00FF00FF00FF00ULL) >> 8;
  x = (x & 0x0F0F0F0F0F0F0F0FULL) << 4 | (x & 0xF0F0F0F0F0F0F0F0ULL) >> 4;
  return x;
}

/* learn/base/localizer.h:106-128  the three key-transform branches (I = uint64). */
static uint64_t transform_key(uint64_t id, uint64_t max_key) {
  if (max_key < UINT64_MAX) return id % max_key; /* :108-115 */
  return orc_reverse_bytes(id);                  /* :116-121 (sizeof(I)==8) */
}

typedef struct { uint64_t k; uint32_t i; } pair_t; /* localizer.h:87-92 */

static int pair_cmp(const void* a, const void* b) {
  uint64_t ka = ((const pair_t*)a)->k, kb = ((const pair_t*)b)->k;
  return ka < kb ? -1 : (ka > kb ? 1 : 0);
}

typedef struct {
  size_t rows, nnz, nuniq;
  int has_value, has_label;
  uint64_t* uniq;   /* ascending transformed keys */
  uint32_t* cnt;    /* multiplicities */
  size_t* offset;   /* rows+1 */
  uint32_t* index;  /* nnz local ids */
  float* value;
  float* label;
} orc_loc_t;

/* learn/base/localizer.h:96-167 CountUniqIndex + :169-221 RemapIndex.
 * dict==NULL: Localize (:42-51), the dictionary is the block's own unique list.
 * dict!=NULL: entries whose key is not in the (ascending) dictionary are dropped. */
static orc_loc_t* localize_impl(const uint64_t* index, const size_t* offset, const float* value,
                                const float* label, size_t rows, uint64_t max_key,
                                const uint64_t* dict, size_t ndict) {
  orc_loc_t* r = (orc_loc_t*)calloc(1, sizeof(orc_loc_t));
  if (rows == 0) return r; /* :101,173 */
  size_t n = offset[rows];
  pair_t* p = (pair_t*)malloc((n ? n : 1) * sizeof(pair_t));
  for (size_t i = 0; i < n; ++i) { p[i].k = transform_key(index[i], max_key); p[i].i = (uint32_t)i; }
  qsort(p, n, sizeof(pair_t), pair_cmp); /* :130-131, tie order does not reach any output */

  /* run-length encode, :139-166 */
  uint64_t* uq = (uint64_t*)malloc((n ? n : 1) * 8);
  uint32_t* ct = (uint32_t*)malloc((n ? n : 1) * 4);
  size_t U = 0;
  for (size_t i = 0; i < n;) {
    size_t j = i;
    while (j < n && p[j].k == p[i].k) ++j;
    uq[U] = p[i].k; ct[U] = (uint32_t)(j - i); ++U; i = j;
  }
  r->uniq = uq; r->cnt = ct; r->nuniq = U;

  const uint64_t* d = dict ? dict : uq;
  size_t nd = dict ? ndict : U;
  if (nd == 0 || n == 0) { /* :173 (empty dict) -- nothing is written */
    if (n == 0 && nd != 0) {
      r->rows = rows; r->offset = (size_t*)calloc(rows + 1, sizeof(size_t));
      if (label) { r->label = (float*)malloc(rows * 4); memcpy(r->label, label, rows * 4); r->has_label = 1; }
    }
    free(p); return r;
  }
  /* two-pointer merge, :179-194: remapped[pair.i] = rank+1 */
  uint32_t* remapped = (uint32_t*)calloc(n, 4);
  size_t cd = 0, cp = 0, matched = 0;
  while (cd < nd && cp < n) {
    if (d[cd] < p[cp].k) { ++cd; }
    else {
      if (d[cd] == p[cp].k) { remapped[p[cp].i] = (uint32_t)(cd + 1); ++matched; }
      ++cp;
    }
  }
  /* rebuild CSR, :199-220 */
  r->rows = rows; r->nnz = matched;
  r->offset = (size_t*)malloc((rows + 1) * sizeof(size_t)); r->offset[0] = 0;
  r->index = (uint32_t*)malloc((matched ? matched : 1) * 4);
  if (value) { r->value = (float*)malloc((matched ? matched : 1) * 4); r->has_value = 1; }
  size_t k = 0;
  for (size_t i = 0; i < rows; ++i) {
    size_t m = 0;
    for (size_t j = offset[i]; j < offset[i + 1]; ++j) {
      if (remapped[j] == 0) continue;
      ++m;
      if (value) r->value[k] = value[j];
      r->index[k++] = remapped[j] - 1;
    }
    r->offset[i + 1] = r->offset[i] + m;
  }
  if (label) { r->label = (float*)malloc(rows * 4); memcpy(r->label, label, rows * 4); r->has_label = 1; }
  free(remapped); free(p);
  return r;
}

ORC_API void* orc_localize(const uint64_t* index, const size_t* offset, const float* value,
                           const float* label, size_t rows, uint64_t max_key) {
  return localize_impl(index, offset, value, label, rows, max_key, NULL, 0);
}
ORC_API void* orc_remap(const uint64_t* index, const size_t* offset, const float* value,
                        const float* label, size_t rows, uint64_t max_key, const uint64_t* dict,
                        size_t ndict) {
  return localize_impl(index, offset, value, label, rows, max_key, dict, ndict);
}
ORC_API size_t orc_loc_nuniq(void* h) { return ((orc_loc_t*)h)->nuniq; }
ORC_API size_t orc_loc_nnz(void* h) { return ((orc_loc_t*)h)->nnz; }
ORC_API size_t orc_loc_rows(void* h) { return ((orc_loc_t*)h)->rows; }
ORC_API int orc_loc_has_value(void* h) { return ((orc_loc_t*)h)->has_value && ((orc_loc_t*)h)->nnz > 0; }
ORC_API int orc_loc_has_label(void* h) { return ((orc_loc_t*)h)->has_label; }
/* cnt_u: counts as unsigned (C=unsigned); cnt_f: cast to float (C=float), localizer.h:139-166 */
ORC_API void orc_loc_get(void* h, uint64_t* uniq, uint32_t* cnt_u, float* cnt_f, size_t* offset,
                         uint32_t* index, float* value, float* label) {
  orc_loc_t* r = (orc_loc_t*)h;
  if (uniq) memcpy(uniq, r->uniq, r->nuniq * 8);
  if (cnt_u) memcpy(cnt_u, r->cnt, r->nuniq * 4);
  if (cnt_f) for (size_t i = 0; i < r->nuniq; ++i) cnt_f[i] = (float)r->cnt[i];
  if (offset && r->offset) memcpy(offset, r->offset, (r->rows + 1) * sizeof(size_t));
  if (index && r->index) memcpy(index, r->index, r->nnz * 4);
  if (value && r->value) memcpy(value, r->value, r->nnz * 4);
  if (label && r->label) memcpy(label, r->label, r->rows * 4);
}
ORC_API void orc_loc_free(void* h) {
  orc_loc_t* r = (orc_loc_t*)h;
  free(r->uniq); free(r->cnt); free(r->offset); free(r->index); free(r->value); free(r->label); free(r);
}

/* ------------------------------------------------------------ SpMV / SpMM -- */

/* learn/base/spmv.h:72-91  y = D x; sequential fp32 sum in CSR order; empty rows untouched (:79). */
ORC_API void orc_spmv_times(size_t rows, const size_t* offset, const uint32_t* index,
                            const float* value, const float* x, float* y) {
  for (size_t i = 0; i < rows; ++i) {
    if (offset[i] == offset[i + 1]) continue;
    float yi = 0;
    if (value) for (size_t j = offset[i]; j < offset[i + 1]; ++j) yi += x[index[j]] * value[j];
    else for (size_t j = offset[i]; j < offset[i + 1]; ++j) yi += x[index[j]];
    y[i] = yi;
  }
}

/* learn/base/spmv.h:95-119  y = D^T x; y zeroed, rows visited in ascending order, so each y[k]
 * is a sequential fp32 sum over ascending rows (the column split over threads does not change it). */
ORC_API void orc_spmv_trans_times(size_t rows, const size_t* offset, const uint32_t* index,
                                  const float* value, const float* x, float* y, size_t ysize) {
  memset(y, 0, ysize * sizeof(float));
  for (size_t i = 0; i < rows; ++i) {
    if (offset[i] == offset[i + 1]) continue;
    float xi = x[i];
    if (value) for (size_t j = offset[i]; j < offset[i + 1]; ++j) y[index[j]] += xi * value[j];
    else for (size_t j = offset[i]; j < offset[i + 1]; ++j) y[index[j]] += xi;
  }
}

/* learn/base/spmm.h:56-81  Y = D X, X row-major with width dim; Y zeroed first (:58). */
ORC_API void orc_spmm_times(size_t rows, const size_t* offset, const uint32_t* index,
                            const float* value, const float* x, float* y, int dim) {
  memset(y, 0, rows * dim * sizeof(float));
  for (size_t i = 0; i < rows; ++i) {
    float* yi = y + i * dim;
    for (size_t j = offset[i]; j < offset[i + 1]; ++j) {
      const float* xj = x + (size_t)index[j] * dim;
      if (value) { float v = value[j]; for (int k = 0; k < dim; ++k) yi[k] += xj[k] * v; }
      else for (int k = 0; k < dim; ++k) yi[k] += xj[k];
    }
  }
}

/* learn/base/spmm.h:85-123  Y = p Z (or 0) then Y[e,:] += X[i,:] * val, rows ascending. */
ORC_API void orc_spmm_trans_times(size_t rows, const size_t* offset, const uint32_t* index,
                                  const float* value, const float* x, float p, const float* z,
                                  float* y, size_t ysize, int dim) {
  if (z && p != 0) { for (size_t i = 0; i < ysize; ++i) y[i] = z[i] * p; } /* spmm.h:47-51,90-94 */
  else memset(y, 0, ysize * sizeof(float));
  for (size_t i = 0; i < rows; ++i) {
    const float* xi = x + i * dim;
    for (size_t j = offset[i]; j < offset[i + 1]; ++j) {
      float* yj = y + (size_t)index[j] * dim;
      if (value) { float v = value[j]; for (int k = 0; k < dim; ++k) yj[k] += xi[k] * v; }
      else for (int k = 0; k < dim; ++k) yj[k] += xi[k];
    }
  }
}

/* ------------------------------------------------------------- evaluation -- */

/* The reference's "#pragma omp parallel for reduction(+:x) num_threads(nt)" with libgomp's
 * static schedule: thread t sums a contiguous chunk into a private 0-initialised copy, the
 * copies are then added to the original. Deterministic for nt <= 2 (a+b == b+a). */
static void omp_static_chunk(size_t n, int nt, int t, size_t* b, size_t* e) {
  size_t q = n / nt, r = n % nt;
  if ((size_t)t < r) { *b = t * (q + 1); *e = *b + q + 1; }
  else { *b = r * (q + 1) + (t - r) * q; *e = *b + q; }
}

typedef struct { float label, predict; } auc_entry_t;
static void auc_merge_sort(auc_entry_t* a, auc_entry_t* tmp, size_t n) {
  if (n < 2) return;
  size_t h = n / 2;
  auc_merge_sort(a, tmp, h); auc_merge_sort(a + h, tmp, n - h);
  size_t i = 0, j = h, k = 0;
  while (i < h && j < n) tmp[k++] = (a[j].predict < a[i].predict) ? a[j++] : a[i++];
  while (i < h) tmp[k++] = a[i++];
  while (j < n) tmp[k++] = a[j++];
  memcpy(a, tmp, n * sizeof(auc_entry_t));
}

/* learn/base/binary_class_evaluation.h:17-38. Float accumulators as in the reference.
 * Entries with equal scores: the reference's std::sort leaves them in an unspecified order;
 * this port keeps input order (stable). Results coincide whenever scores are distinct. */
ORC_API float orc_auc(const float* label, const float* predict, size_t n) {
  auc_entry_t* buf = (auc_entry_t*)malloc((n ? n : 1) * sizeof(auc_entry_t));
  auc_entry_t* tmp = (auc_entry_t*)malloc((n ? n : 1) * sizeof(auc_entry_t));
  for (size_t i = 0; i < n; ++i) { buf[i].label = label[i]; buf[i].predict = predict[i]; }
  auc_merge_sort(buf, tmp, n);
  float area = 0, cum_tp = 0;
  for (size_t i = 0; i < n; ++i) { if (buf[i].label > 0) cum_tp += 1; else area += cum_tp; }
  free(buf); free(tmp);
  if (cum_tp == 0 || cum_tp == n) return 1;
  area /= cum_tp * (n - cum_tp);
  return area < 0.5 ? 1 - area : area;
}

/* binary_class_evaluation.h:40-51 */
ORC_API float orc_accuracy(const float* label, const float* predict, size_t n, float threshold) {
  float correct = 0;
  for (size_t i = 0; i < n; ++i)
    if ((label[i] > 0 && predict[i] > threshold) || (label[i] <= 0 && predict[i] <= threshold)) correct += 1;
  float acc = correct / (float)n;
  return acc > 0.5 ? acc : 1 - acc;
}

/* binary_class_evaluation.h:66-74 */
ORC_API float orc_logit_objv(const float* label, const float* predict, size_t n, int nt) {
  float objv = 0;
  for (int t = 0; t < nt; ++t) {
    size_t b, e; omp_static_chunk(n, nt, t, &b, &e);
    float part = 0;
    for (size_t i = b; i < e; ++i) {
      float y = label[i] > 0 ? 1 : -1;
      part += logf(1 + expf(-y * predict[i]));
    }
    objv += part;
  }
  return objv;
}

/* binary_class_evaluation.h:53-64 */
ORC_API float orc_logloss(const float* label, const float* predict, size_t n, int nt) {
  float loss = 0;
  for (int t = 0; t < nt; ++t) {
    size_t b, e; omp_static_chunk(n, nt, t, &b, &e);
    float part = 0;
    for (size_t i = b; i < e; ++i) {
      float y = label[i] > 0;
      float p = 1 / (1 + expf(-predict[i]));
      p = p < 1e-10 ? 1e-10 : p;
      part += y * logf(p) + (1 - y) * logf(1 - p);
    }
    loss += part;
  }
  return -loss;
}

/* -------------------------------------------------------------- the loss -- */

/* learn/linear/loss.h. loss_type: 2 = LOGIT, 4 = SQUARE_HINGE (config.proto:30-38).
 * Init (:32-39) Xw = X w; Evaluate (:42-46,81-86,100-104,128-139) fills
 * prog = {objv, acc, auc, count, new_ex, new_w}; CalcGrad (:106-115,141-156) if grad != NULL.
 * xw (rows) and dual (rows) are optional outputs for inspection. */
ORC_API void orc_linear_loss(int loss_type, size_t rows, const size_t* offset, const uint32_t* index,
                             const float* value, const float* label, const float* w, size_t U, int nt,
                             double* prog, float* grad, float* xw_out, float* dual_out) {
  float* xw = (float*)calloc(rows ? rows : 1, sizeof(float));
  orc_spmv_times(rows, offset, index, value, w, xw);
  for (int i = 0; i < 6; ++i) prog[i] = 0;
  prog[4] = (double)rows; prog[3] = 1;
  prog[2] = orc_auc(label, xw, rows);
  prog[1] = orc_accuracy(label, xw, rows, 0);
  if (loss_type == 2) {
    prog[0] = orc_logit_objv(label, xw, rows, nt);
  } else {
    float objv = 0;
    for (int t = 0; t < nt; ++t) {
      size_t b, e; omp_static_chunk(rows, nt, t, &b, &e);
      float part = 0;
      for (size_t i = b; i < e; ++i) {
        float y = label[i] > 0 ? 1 : -1;
        float tmp = 1 - y * xw[i]; if (tmp < 0) tmp = 0;
        part += tmp * tmp;
      }
      objv += part;
    }
    prog[0] = objv;
  }
  if (xw_out) memcpy(xw_out, xw, rows * sizeof(float));
  if (grad || dual_out) {
    float* dual = (float*)malloc((rows ? rows : 1) * sizeof(float));
    for (size_t i = 0; i < rows; ++i) {
      float y = label[i] > 0 ? 1 : -1;
      if (loss_type == 2) dual[i] = -y / (1 + expf(y * xw[i]));   /* :112 */
      else dual[i] = y * (y * xw[i] > 1.0);                       /* :148, as written */
    }
    if (dual_out) memcpy(dual_out, dual, rows * sizeof(float));
    if (grad) {
      orc_spmv_trans_times(rows, offset, index, value, dual, grad, U);
      if (loss_type != 2) for (size_t i = 0; i < U; ++i) grad[i] *= -2.0; /* :152-155 (double multiply, float store) */
    }
    free(dual);
  }
  free(xw);
}

/* ----------------------------------------------------------- the updaters -- */

/* learn/linear/penalty.h:36-41 */
ORC_API float orc_l1l2_solve(float l1, float l2, float z, float eta) {
  if (z <= l1 && z >= -l1) return 0;
  return (z > 0 ? z - l1 : z + l1) / (eta + l2);
}

/* One push message applied to caller-held entries. algo: 1 SGD, 2 ADAGRAD, 3 FTRL
 * (config.proto:45-52). Defaults alpha=.1 beta=1 are overridden only by present conf keys
 * (learn/linear/async_sgd.h:49,205-206). Returns the change of |w|_0 (:35-41). */
ORC_API int64_t orc_handle_push(int algo, float l1, float l2, int has_eta, float eta_c, int has_beta,
                                float beta_c, size_t n, const float* grad, float* w, float* z,
                                float* cg, int* t_inout) {
  float alpha = has_eta ? eta_c : 0.1f, beta = has_beta ? beta_c : 1.0f;
  int64_t new_w = 0;
  float eta_sgd = 0;
  if (algo == 1) { /* SGDHandle::Start, async_sgd.h:85-90 */
    eta_sgd = (beta + sqrtf((float)*t_inout)) / alpha;
    *t_inout += 1;
  }
  for (size_t i = 0; i < n; ++i) {
    float g = grad[i], old_w = w[i], nw;
    if (algo == 1) {                      /* SGDHandle::Push :91-95 */
      nw = orc_l1l2_solve(l1, l2, eta_sgd * old_w - g, eta_sgd);
    } else if (algo == 2) {               /* AdaGradHandle::Push :123-135 */
      float sqrt_n = cg[i];
      cg[i] = sqrtf(sqrt_n * sqrt_n + g * g);
      float eta = (cg[i] + beta) / alpha;
      nw = orc_l1l2_solve(l1, l2, eta * old_w - g, eta);
    } else {                              /* FTRLHandle::Push :160-175 */
      float sqrt_n = cg[i];
      cg[i] = sqrtf(sqrt_n * sqrt_n + g * g);
      float sigma = (cg[i] - sqrt_n) / alpha;
      z[i] += g - sigma * old_w;
      nw = orc_l1l2_solve(l1, l2, -z[i], (beta + cg[i]) / alpha);
    }
    w[i] = nw;
    if (old_w == 0 && nw != 0) ++new_w; else if (old_w != 0 && nw == 0) --new_w; /* :35-41 */
  }
  return new_w;
}

/* ------------------------------------------------ full linear job (1w/1s) -- */

typedef struct {
  uint64_t* keys; float *w, *z, *cg; uint8_t* used; size_t cap, size;
} orc_store_t;

static uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33; return x;
}
static void store_init(orc_store_t* s, size_t cap) {
  s->cap = cap; s->size = 0;
  s->keys = (uint64_t*)calloc(cap, 8); s->w = (float*)calloc(cap, 4); s->z = (float*)calloc(cap, 4);
  s->cg = (float*)calloc(cap, 4); s->used = (uint8_t*)calloc(cap, 1);
}
static void store_free(orc_store_t* s) { free(s->keys); free(s->w); free(s->z); free(s->cg); free(s->used); }
static size_t store_slot(orc_store_t* s, uint64_t key);
static void store_grow(orc_store_t* s) {
  orc_store_t n; store_init(&n, s->cap * 2);
  for (size_t i = 0; i < s->cap; ++i) if (s->used[i]) {
    size_t j = store_slot(&n, s->keys[i]);
    n.w[j] = s->w[i]; n.z[j] = s->z[i]; n.cg[j] = s->cg[i];
  }
  store_free(s); *s = n;
}
/* entries are default-constructed (all zero) on first touch, as std::unordered_map::operator[] */
static size_t store_slot(orc_store_t* s, uint64_t key) {
  if ((s->size + 1) * 2 > s->cap) store_grow(s);
  size_t m = s->cap - 1, j = mix64(key) & m;
  while (s->used[j] && s->keys[j] != key) j = (j + 1) & m;
  if (!s->used[j]) { s->used[j] = 1; s->keys[j] = key; s->size++; }
  return j;
}

typedef struct {
  int algo, loss, has_eta, has_beta, t;
  float l1, l2, eta, beta;
  uint64_t max_key;
  orc_store_t store;
} orc_job_t;

ORC_API void* orc_linear_create(int algo, int loss, float l1, float l2, int has_eta, float eta,
                                int has_beta, float beta, uint64_t max_key) {
  orc_job_t* j = (orc_job_t*)calloc(1, sizeof(orc_job_t));
  j->algo = algo; j->loss = loss; j->l1 = l1; j->l2 = l2; j->has_eta = has_eta; j->eta = eta;
  j->has_beta = has_beta; j->beta = beta; j->max_key = max_key; j->t = 1; /* SGDHandle::t = 1, :100 */
  store_init(&j->store, 1 << 16);
  return j;
}

/* learn/linear/async_sgd.h:240-288 AsgdWorker::ProcessMinibatch with 1 worker, 1 server:
 * Localize -> ZPull (Handle::Pull per key, ascending) -> Init/Evaluate -> CalcGrad -> ZPush
 * (Handle::Start + Push per key + Finish). nt is the worker's hard-coded 2 (:303).
 * prog = {objv, acc, auc, count, new_ex, new_w}. */
ORC_API void orc_linear_minibatch(void* h, const uint64_t* index, const size_t* offset,
                                  const float* value, const float* label, size_t rows, int train,
                                  double* prog) {
  orc_job_t* j = (orc_job_t*)h;
  for (int i = 0; i < 6; ++i) prog[i] = 0;
  if (rows == 0) return;
  orc_loc_t* d = localize_impl(index, offset, value, label, rows, j->max_key, NULL, 0);
  size_t U = d->nuniq;
  float* val = (float*)malloc((U ? U : 1) * 4);
  size_t* slot = (size_t*)malloc((U ? U : 1) * sizeof(size_t));
  for (size_t u = 0; u < U; ++u) store_slot(&j->store, d->uniq[u]);       /* create first: growth moves slots */
  for (size_t u = 0; u < U; ++u) { slot[u] = store_slot(&j->store, d->uniq[u]); val[u] = j->store.w[slot[u]]; }
  float* grad = train ? (float*)malloc((U ? U : 1) * 4) : NULL;
  orc_linear_loss(j->loss, rows, d->offset, d->index, d->value, d->label, val, U, 2, prog, grad, NULL, NULL);
  if (train) {
    float *w = (float*)malloc((U ? U : 1) * 4), *z = (float*)malloc((U ? U : 1) * 4), *cg = (float*)malloc((U ? U : 1) * 4);
    for (size_t u = 0; u < U; ++u) { w[u] = j->store.w[slot[u]]; z[u] = j->store.z[slot[u]]; cg[u] = j->store.cg[slot[u]]; }
    prog[5] = (double)orc_handle_push(j->algo, j->l1, j->l2, j->has_eta, j->eta, j->has_beta, j->beta,
                                      U, grad, w, z, cg, &j->t);
    for (size_t u = 0; u < U; ++u) { j->store.w[slot[u]] = w[u]; j->store.z[slot[u]] = z[u]; j->store.cg[slot[u]] = cg[u]; }
    free(w); free(z); free(cg); free(grad);
  }
  free(val); free(slot);
  orc_loc_free(d);
}

ORC_API size_t orc_linear_model_size(void* h) { return ((orc_job_t*)h)->store.size; }

typedef struct { uint64_t k; size_t slot; } ks_t;
static int ks_cmp(const void* a, const void* b) {
  uint64_t ka = ((const ks_t*)a)->k, kb = ((const ks_t*)b)->k; return ka < kb ? -1 : (ka > kb);
}
/* keys ascending */
ORC_API void orc_linear_model_get(void* h, uint64_t* keys, float* w, float* z, float* cg) {
  orc_store_t* s = &((orc_job_t*)h)->store;
  ks_t* a = (ks_t*)malloc((s->size ? s->size : 1) * sizeof(ks_t));
  size_t n = 0;
  for (size_t i = 0; i < s->cap; ++i) if (s->used[i]) { a[n].k = s->keys[i]; a[n].slot = i; ++n; }
  qsort(a, n, sizeof(ks_t), ks_cmp);
  for (size_t i = 0; i < n; ++i) {
    keys[i] = a[i].k; w[i] = s->w[a[i].slot];
    if (z) z[i] = s->z[a[i].slot];
    if (cg) cg[i] = s->cg[a[i].slot];
  }
  free(a);
}
ORC_API void orc_linear_destroy(void* h) { store_free(&((orc_job_t*)h)->store); free(h); }
