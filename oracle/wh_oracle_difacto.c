/*
 * wh_oracle_difacto.c -- CPU restatement ("port") of the DiFacto part of the hot path:
 * the factorization-machine loss (learn/difacto/loss.h) and the server's update handle
 * (learn/difacto/async_sgd.h), in plain C.
 *
 * TEST INFRASTRUCTURE ONLY (see wh_oracle.c). Parity status: PINNED -- checked bit for bit
 * against oracle/_ref/libwh_ref_difacto.so (the reference's own headers) by
 * tests/test_oracle_vs_ref.py and against tests/golden/golden_difacto.npz. Unpinned, as for
 * the linear part: the ps-lite v1 server loop (restated in oracle/shims/ps.h), and which libm
 * overloads ps-lite's headers bring into scope for learn/difacto (float ones assumed).
 * V is initialised with glibc rand() in server arrival order (async_sgd.h:254): this port
 * calls the same rand(), so it matches the reference for the same srand() seed.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* from wh_oracle.c */
void* orc_localize(const uint64_t*, const size_t*, const float*, const float*, size_t, uint64_t);
size_t orc_loc_nuniq(void*);
size_t orc_loc_nnz(void*);
void orc_loc_get(void*, uint64_t*, uint32_t*, float*, size_t*, uint32_t*, float*, float*);
void orc_loc_free(void*);
float orc_auc(const float*, const float*, size_t);
float orc_logit_objv(const float*, const float*, size_t, int);

/* ------------------------------------------------------------------ the loss ------------ */

/* learn/difacto/loss.h:24-41 (ctor), :185-253 (Data::Load), :53-94 (Evaluate), :102-158 (CalcGrad),
 * :256-264 (Data::Save). model = concatenation over keys of model_siz[k] floats, model_siz in
 * {0, 1, dim+1}. prog = {objv, auc, objv_w, copc, count, new_ex, new_w, new_V} (progress.h:28-36).
 * grad (same layout as model) may be NULL. dropout > 0 draws from this process's rand() (loss.h:150-154), one value per
 * embedding-gradient float in key order: callers srand() first. */
ORC_API void orc_difacto_loss(size_t rows, const size_t* offset, const uint32_t* index, const float* value,
                              const float* label, const float* model, size_t model_len, const int* model_siz,
                              size_t nkeys, int dim, int nt, float clip, float dropout, float norm, double* prog,
                              float* grad) {
  (void)model_len;
  /* w.Load(0, ...): :191-203 */
  int64_t* posw = (int64_t*)malloc((nkeys ? nkeys : 1) * sizeof(int64_t));
  float* w = (float*)calloc(nkeys ? nkeys : 1, sizeof(float));
  size_t p = 0;
  for (size_t i = 0; i < nkeys; ++i) {
    if (model_siz[i] == 0) posw[i] = -1;
    else { posw[i] = (int64_t)p; w[i] = model[p]; p += model_siz[i]; }
  }
  /* V.Load(dim, ...): :204-219 -- the columns whose entry carries an embedding */
  uint32_t* col_map = (uint32_t*)calloc(nkeys ? nkeys : 1, sizeof(uint32_t));
  size_t* posv = (size_t*)malloc((nkeys ? nkeys : 1) * sizeof(size_t));
  size_t m = 0;
  if (dim > 0) {
    p = 0;
    for (size_t i = 0; i < nkeys; ++i) {
      if (model_siz[i] == dim + 1) { posv[m] = p + 1; col_map[i] = (uint32_t)(++m); }
      p += model_siz[i];
    }
  }
  float* V = (float*)malloc(((m * dim) ? m * dim : 1) * sizeof(float));
  for (size_t i = 0; i < m; ++i) memcpy(V + i * dim, model + posv[i], dim * sizeof(float));
  /* X restricted to those columns: :224-241; XX = X with squared values: :245-252 */
  size_t n = rows ? offset[rows] : 0;
  size_t* os = (size_t*)malloc((rows + 1) * sizeof(size_t));
  uint32_t* vidx = (uint32_t*)malloc((n ? n : 1) * sizeof(uint32_t));
  float* vval = value ? (float*)malloc((n ? n : 1) * sizeof(float)) : NULL;
  float* vval2 = value ? (float*)malloc((n ? n : 1) * sizeof(float)) : NULL;
  size_t k = 0;
  os[0] = 0;
  if (m) {
    for (size_t i = 0; i < rows; ++i) {
      for (size_t j = offset[i]; j < offset[i + 1]; ++j) {
        uint32_t c = col_map[index[j]];
        if (c > 0) { vidx[k] = c - 1; if (value) { vval[k] = value[j]; vval2[k] = value[j] * value[j]; } ++k; }
      }
      os[i + 1] = k;
    }
  }

  /* Evaluate: py = X w (:56-58) */
  float* py = (float*)calloc(rows ? rows : 1, sizeof(float));
  for (size_t i = 0; i < rows; ++i) {
    if (offset[i] == offset[i + 1]) continue;
    float y = 0;
    for (size_t j = offset[i]; j < offset[i + 1]; ++j) y += value ? w[index[j]] * value[j] : w[index[j]];
    py[i] = y;
  }
  for (int i = 0; i < 8; ++i) prog[i] = 0;
  prog[2] = orc_logit_objv(label, py, rows, nt); /* :60-61 */
  float* XV = NULL;
  if (m) {
    /* :64-84 */
    float* xxvv = (float*)calloc((rows * dim) ? rows * dim : 1, sizeof(float));
    XV = (float*)calloc((rows * dim) ? rows * dim : 1, sizeof(float));
    for (size_t i = 0; i < rows; ++i) {
      for (size_t j = os[i]; j < os[i + 1]; ++j) {
        const float* vr = V + (size_t)vidx[j] * dim;
        for (int d = 0; d < dim; ++d) {
          float vv = vr[d] * vr[d];
          if (value) { xxvv[i * dim + d] += vv * vval2[j]; XV[i * dim + d] += vr[d] * vval[j]; }
          else { xxvv[i * dim + d] += vv; XV[i * dim + d] += vr[d]; }
        }
      }
    }
    for (size_t i = 0; i < rows; ++i) {
      float s = 0;
      for (int d = 0; d < dim; ++d) s += XV[i * dim + d] * XV[i * dim + d] - xxvv[i * dim + d];
      py[i] += .5 * s; /* double product, float store: :82 */
    }
    free(xxvv);
    prog[0] = orc_logit_objv(label, py, rows, nt);
  } else {
    prog[0] = prog[2];
  }
  prog[1] = orc_auc(label, py, rows); /* :90 */
  prog[5] = (double)rows; prog[4] = 1;

  if (grad) {
    /* CalcGrad: :105-112 */
    for (size_t i = 0; i < rows; ++i) {
      float y = label[i] > 0 ? 1 : -1;
      py[i] = -y / (1 + expf(y * py[i]));
    }
    float* gw = (float*)calloc(nkeys ? nkeys : 1, sizeof(float));
    for (size_t i = 0; i < rows; ++i)
      for (size_t j = offset[i]; j < offset[i + 1]; ++j) gw[index[j]] += value ? py[i] * value[j] : py[i];
    for (size_t i = 0; i < nkeys; ++i) if (posw[i] >= 0) grad[posw[i]] = gw[i]; /* w.Save: :256-264 */
    free(gw);
    if (m) {
      /* xxp = (X.*X)' p : :121-123 */
      float* xxp = (float*)calloc(m, sizeof(float));
      for (size_t i = 0; i < rows; ++i)
        for (size_t j = os[i]; j < os[i + 1]; ++j) xxp[vidx[j]] += value ? py[i] * vval2[j] : py[i];
      /* V = -diag(xxp) V : :126-131 */
      for (size_t i = 0; i < m; ++i) for (int d = 0; d < dim; ++d) V[i * dim + d] *= -xxp[i];
      /* XV = diag(p) X V : :134-140 */
      for (size_t i = 0; i < rows; ++i) for (int d = 0; d < dim; ++d) XV[i * dim + d] *= py[i];
      /* V += X' XV : :142 (y = z * 1 in place, then the accumulation in row order) */
      for (size_t i = 0; i < m * dim; ++i) V[i] = V[i] * 1.0f;
      for (size_t i = 0; i < rows; ++i)
        for (size_t j = os[i]; j < os[i + 1]; ++j) {
          float* yr = V + (size_t)vidx[j] * dim;
          for (int d = 0; d < dim; ++d) yr[d] += value ? XV[i * dim + d] * vval[j] : XV[i * dim + d];
        }
      if (clip > 0) for (size_t i = 0; i < m * dim; ++i) V[i] = V[i] > clip ? clip : (V[i] < -clip ? -clip : V[i]); /* :145-148 */
      if (dropout > 0) /* :150-154 */
        for (size_t i = 0; i < m * dim; ++i) if ((float)rand() / RAND_MAX > 1 - dropout) V[i] = 0;
      if (norm) { /* :155,160-166 */
        float nm = 0;
        for (size_t i = 0; i < m * dim; ++i) nm += V[i] * V[i];
        if (!(nm < 1e-10)) { nm = sqrtf(nm); for (size_t i = 0; i < m * dim; ++i) V[i] = V[i] / nm; }
      }
      for (size_t i = 0; i < m; ++i) memcpy(grad + posv[i], V + i * dim, dim * sizeof(float)); /* V.Save */
      free(xxp);
    }
  }
  free(posw); free(w); free(col_map); free(posv); free(V); free(os); free(vidx); free(vval); free(vval2);
  free(py); free(XV);
}

/* ------------------------------------------------------------------ the server ---------- */

typedef struct {
  float l1, l2, alpha, beta;       /* for w */
  int dim; unsigned thr; float v_l2, v_alpha, v_beta, v_min, v_max; /* for V */
  int l1_shrk;
  int64_t new_w, new_V;
} dfc_handle_t;

typedef struct {
  uint64_t* keys; uint8_t* used; size_t cap, size;
  uint32_t* fea_cnt; int* esize;
  float* w;   /* cap * (dim+1): {w0, V[dim]} */
  float* cg;  /* cap * (dim+2): {sqc_grad_0, z_0, sqc_grad_V[dim]} */
  int dim;
} dfc_store_t;

static uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33; return x;
}
static void dstore_init(dfc_store_t* s, size_t cap, int dim) {
  s->cap = cap; s->size = 0; s->dim = dim;
  s->keys = (uint64_t*)calloc(cap, 8); s->used = (uint8_t*)calloc(cap, 1);
  s->fea_cnt = (uint32_t*)calloc(cap, 4); s->esize = (int*)calloc(cap, sizeof(int));
  s->w = (float*)calloc(cap * (dim + 1), 4); s->cg = (float*)calloc(cap * (dim + 2), 4);
}
static void dstore_free(dfc_store_t* s) { free(s->keys); free(s->used); free(s->fea_cnt); free(s->esize); free(s->w); free(s->cg); }
static size_t dstore_slot(dfc_store_t* s, uint64_t key);
static void dstore_grow(dfc_store_t* s) {
  dfc_store_t n; dstore_init(&n, s->cap * 2, s->dim);
  for (size_t i = 0; i < s->cap; ++i) if (s->used[i]) {
    size_t j = dstore_slot(&n, s->keys[i]);
    n.fea_cnt[j] = s->fea_cnt[i]; n.esize[j] = s->esize[i];
    memcpy(n.w + j * (s->dim + 1), s->w + i * (s->dim + 1), (s->dim + 1) * 4);
    memcpy(n.cg + j * (s->dim + 2), s->cg + i * (s->dim + 2), (s->dim + 2) * 4);
  }
  dstore_free(s); *s = n;
}
/* AdaGradEntry default: fea_cnt 0, size 1, all zero (async_sgd.h:196-208) */
static size_t dstore_slot(dfc_store_t* s, uint64_t key) {
  if ((s->size + 1) * 2 > s->cap) dstore_grow(s);
  size_t mk = s->cap - 1, j = mix64(key) & mk;
  while (s->used[j] && s->keys[j] != key) j = (j + 1) & mk;
  if (!s->used[j]) { s->used[j] = 1; s->keys[j] = key; s->esize[j] = 1; s->size++; }
  return j;
}

/* AdaGradHandle::Resize, async_sgd.h:247-259 */
static void dfc_resize(dfc_handle_t* h, dfc_store_t* s, size_t j) {
  float* w = s->w + j * (s->dim + 1); float* cg = s->cg + j * (s->dim + 2);
  if (s->fea_cnt[j] > h->thr && s->esize[j] < h->dim + 1 && (!h->l1_shrk || w[0] != 0)) {
    int old = s->esize[j];
    s->esize[j] = h->dim + 1;
    for (int k = old; k < s->esize[j]; ++k) {
      w[k] = rand() / (float)RAND_MAX * (h->v_max - h->v_min) + h->v_min;
      cg[k + 1] = 0;
    }
    h->new_V += s->esize[j] - old;
  }
}
/* AdaGradHandle::UpdateW, async_sgd.h:262-286 */
static void dfc_update_w(dfc_handle_t* h, dfc_store_t* s, size_t j, float g) {
  float* wv = s->w + j * (s->dim + 1); float* cgv = s->cg + j * (s->dim + 2);
  float w = wv[0];
  g += h->l2 * w;
  float cg = cgv[0];
  float cg_new = sqrtf(cg * cg + g * g);
  cgv[0] = cg_new;
  cgv[1] -= g - (cg_new - cg) / h->alpha * w;
  float z = cgv[1], l1 = h->l1;
  if (z <= l1 && z >= -l1) wv[0] = 0;
  else { float eta = (h->beta + cg_new) / h->alpha; wv[0] = (z > 0 ? z - l1 : z + l1) / eta; }
  if (w == 0 && wv[0] != 0) { ++h->new_w; dfc_resize(h, s, j); }
  else if (w != 0 && wv[0] == 0) { --h->new_w; }
}
/* AdaGradHandle::UpdateV, async_sgd.h:289-296 */
static void dfc_update_v(dfc_handle_t* h, float* w, float* cg, const float* g, int n) {
  for (int i = 0; i < n; ++i) {
    float grad = g[i] + h->v_l2 * w[i];
    cg[i] = sqrtf(cg[i] * cg[i] + grad * grad);
    float eta = h->v_alpha / (cg[i] + h->v_beta);
    w[i] -= eta * grad;
  }
}

typedef struct {
  dfc_handle_t h; dfc_store_t store; uint64_t max_key;
  float clip, dropout, norm; int nt, do_embedding;
} dfc_job_t;

/* AsyncServer ctor, async_sgd.h:302-327 */
ORC_API void* orc_difacto_create(float l1, float l2, float eta, float beta, int l1_shrk, int dim, int threshold,
                                 float v_l2, int v_has_eta, float v_eta, int v_has_beta, float v_beta,
                                 float init_scale, float dropout, float clip, float norm, int nt, uint64_t max_key,
                                 unsigned seed) {
  dfc_job_t* j = (dfc_job_t*)calloc(1, sizeof(dfc_job_t));
  srand(seed);
  j->h.l1 = l1; j->h.l2 = l2; j->h.alpha = eta; j->h.beta = beta; j->h.l1_shrk = l1_shrk;
  j->h.dim = dim; j->h.thr = (unsigned)threshold; j->h.v_l2 = v_l2;
  j->h.v_min = -init_scale; j->h.v_max = init_scale;
  j->h.v_alpha = v_has_eta ? v_eta : eta; j->h.v_beta = v_has_beta ? v_beta : beta;
  if (dim == 0) { j->h.thr = 0; j->h.v_l2 = 0; j->h.v_alpha = .01f; j->h.v_beta = 1; j->h.v_min = -.01f; j->h.v_max = .01f; }
  j->max_key = max_key; j->clip = clip; j->dropout = dropout; j->norm = norm; j->nt = nt; j->do_embedding = dim > 0;
  dstore_init(&j->store, 1 << 16, dim);
  return j;
}

/* AsyncWorker::ProcessMinibatch, async_sgd.h:363-425, 1 worker + 1 server:
 * Localize (float counts) -> [pass 0: count push, cmd = kPushFeaCnt] -> ZVPull -> Loss::Evaluate ->
 * CalcGrad -> ZVPush. prog as in orc_difacto_loss plus new_w / new_V of this minibatch. */
ORC_API void orc_difacto_minibatch(void* hnd, const uint64_t* index, const size_t* offset, const float* value,
                                   const float* label, size_t rows, int train, int data_pass, double* prog) {
  dfc_job_t* j = (dfc_job_t*)hnd;
  dfc_handle_t* h = &j->h; dfc_store_t* s = &j->store;
  int dim = h->dim;
  for (int i = 0; i < 8; ++i) prog[i] = 0;
  if (rows == 0) return;
  void* loc = orc_localize(index, offset, value, label, rows, j->max_key);
  size_t U = orc_loc_nuniq(loc), n = orc_loc_nnz(loc);
  uint64_t* keys = (uint64_t*)malloc((U ? U : 1) * 8);
  float* cnt = (float*)malloc((U ? U : 1) * 4);
  size_t* loff = (size_t*)malloc((rows + 1) * sizeof(size_t));
  uint32_t* lidx = (uint32_t*)malloc((n ? n : 1) * 4);
  float* lval = value ? (float*)malloc((n ? n : 1) * 4) : NULL;
  float* llab = (float*)malloc(rows * 4);
  orc_loc_get(loc, keys, NULL, cnt, loff, lidx, lval, llab);
  orc_loc_free(loc);
  h->new_w = 0; h->new_V = 0;
  for (size_t u = 0; u < U; ++u) dstore_slot(s, keys[u]); /* create first: growth moves slots */
  size_t* slot = (size_t*)malloc((U ? U : 1) * sizeof(size_t));
  for (size_t u = 0; u < U; ++u) slot[u] = dstore_slot(s, keys[u]);
  if (train && data_pass == 0 && j->do_embedding) { /* :374-381; Push with push_count: :217-219 */
    for (size_t u = 0; u < U; ++u) { s->fea_cnt[slot[u]] += (unsigned)cnt[u]; dfc_resize(h, s, slot[u]); }
  }
  /* ZVPull: AdaGradHandle::Pull, :234-244 */
  int* siz = (int*)malloc((U ? U : 1) * sizeof(int));
  size_t len = 0;
  for (size_t u = 0; u < U; ++u) {
    float w0 = s->w[slot[u] * (dim + 1)];
    siz[u] = (s->esize[slot[u]] == 1 || (h->l1_shrk && w0 == 0)) ? 1 : s->esize[slot[u]];
    len += siz[u];
  }
  float* model = (float*)malloc((len ? len : 1) * 4);
  float* grad = (float*)malloc((len ? len : 1) * 4);
  size_t p = 0;
  for (size_t u = 0; u < U; ++u) { memcpy(model + p, s->w + slot[u] * (dim + 1), siz[u] * 4); p += siz[u]; }
  orc_difacto_loss(rows, loff, lidx, lval, llab, model, len, siz, U, dim, j->nt, j->clip, j->dropout, j->norm, prog,
                   train ? grad : NULL);
  if (train) { /* ZVPush: Push without push_count, :220-231 */
    p = 0;
    for (size_t u = 0; u < U; ++u) {
      dfc_update_w(h, s, slot[u], grad[p]);
      if (siz[u] > 1) dfc_update_v(h, s->w + slot[u] * (dim + 1) + 1, s->cg + slot[u] * (dim + 2) + 2, grad + p + 1, siz[u] - 1);
      p += siz[u];
    }
  }
  prog[6] = (double)h->new_w; prog[7] = (double)h->new_V;
  free(keys); free(cnt); free(loff); free(lidx); free(lval); free(llab); free(slot); free(siz); free(model); free(grad);
}

ORC_API size_t orc_difacto_model_size(void* hnd) { return ((dfc_job_t*)hnd)->store.size; }

typedef struct { uint64_t k; size_t slot; } dks_t;
static int dks_cmp(const void* a, const void* b) {
  uint64_t ka = ((const dks_t*)a)->k, kb = ((const dks_t*)b)->k; return ka < kb ? -1 : (ka > kb);
}
/* keys ascending; w rows of dim+1, cg rows of dim+2, zero padded beyond the entry's size */
ORC_API void orc_difacto_model_get(void* hnd, int dim, uint64_t* keys, uint32_t* fea_cnt, int* size, float* w, float* cg) {
  dfc_store_t* s = &((dfc_job_t*)hnd)->store;
  dks_t* a = (dks_t*)malloc((s->size ? s->size : 1) * sizeof(dks_t));
  size_t n = 0;
  for (size_t i = 0; i < s->cap; ++i) if (s->used[i]) { a[n].k = s->keys[i]; a[n].slot = i; ++n; }
  qsort(a, n, sizeof(dks_t), dks_cmp);
  for (size_t i = 0; i < n; ++i) {
    size_t j = a[i].slot;
    keys[i] = a[i].k; fea_cnt[i] = s->fea_cnt[j]; size[i] = s->esize[j];
    for (int k = 0; k < dim + 1; ++k) w[i * (dim + 1) + k] = k < s->esize[j] ? s->w[j * (dim + 1) + k] : 0;
    for (int k = 0; k < dim + 2; ++k) cg[i * (dim + 2) + k] = (k < 2 || k - 1 < s->esize[j]) ? s->cg[j * (dim + 2) + k] : 0;
  }
  free(a);
}
ORC_API void orc_difacto_destroy(void* hnd) { dstore_free(&((dfc_job_t*)hnd)->store); free(hnd); }
