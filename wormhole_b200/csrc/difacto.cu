// Piece (5): DiFacto -- the factorization-machine loss (learn/difacto/loss.h) and the server's
// FTRL-for-w / AdaGrad-for-V handle with lazily created embeddings (learn/difacto/async_sgd.h).
//
// Device layout. The reference moves variable-length entries (1 or 1+dim floats per key) and
// allocates V on the heap when a feature has been seen often enough. Here a shard is dense and
// direct-addressed: base float4 {w0, z0, sqc_grad_0, fea_cnt bits}, a byte "has V", and two
// [keys x dim] float tables for V and its sqrt-cumulative gradient (1 GB each at 2^24 keys, dim 16:
// HBM is large enough that laziness would only add indirection). Between kernels a minibatch's
// model lives in a strided form: w[U], has[U], V[U x dim]; the concatenated ZVPull / ZVPush
// payload of the reference is only built at the C-ABI boundary (wh_difacto_loss).
//
// Forward sums (X w, X V, (X.X)(V.V)) walk each row in CSR order: bit-exact with the reference.
// Transposed sums use the fixed-point accumulators of spmv.cuh.
#include "batch.cuh"
#include "loss.cuh"
#include "scan.cuh"
#include "spmv.cuh"

namespace wh {

int spmv_trans_accumulate_launch(whContext* ctx, size_t rows, size_t nnz, const size_t* offset,
                                 const uint32_t* index, const float* value, const float* x,
                                 long long* acc, FixedScale* fs, unsigned copies, size_t stride, bool scale_ready = false);
int fixed_finish_launch(whContext* ctx, long long* acc, size_t n, const FixedScale* fs, float* y, float mul,
                        unsigned copies, size_t stride, const unsigned* d_n = nullptr, unsigned ystride = 1);

// ------------------------------------------------------------------------------ the loss --

// The column subset of the minibatch for V (Data::Load, learn/difacto/loss.h:204-253: V.X holds only the entries whose
// key came with an embedding): per row the run {vstart, vcnt} of its entries with has[key], in CSR order, as (key rank,
// value) pairs. One thread per row; a row reserves its run with one atomic (where the runs lie does not matter, their
// inner order does). Keys with an embedding are few (fea_cnt > threshold and w != 0), so everything downstream walks a
// few entries per row instead of testing all of them once per embedding dimension.
__global__ void __launch_bounds__(256)
dfc_subset_kernel(size_t rows, const size_t* __restrict__ offset, const uint32_t* __restrict__ index,
                  const float* __restrict__ value, const unsigned char* __restrict__ has, unsigned* cursor,
                  unsigned* __restrict__ vstart, unsigned* __restrict__ vcnt, uint32_t* __restrict__ vidx,
                  float* __restrict__ vval) {
  const size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const size_t b = offset[r], e = offset[r + 1];
  unsigned n = 0;
  for (size_t j = b; j < e; ++j) n += has[index[j]] ? 1u : 0u;
  unsigned at = 0;
  if (n) at = atomicAdd(cursor, n);
  vstart[r] = at; vcnt[r] = n;
  if (!n) return;
  for (size_t j = b; j < e; ++j) {
    const uint32_t c = index[j];
    if (!has[c]) continue;
    vidx[at] = c;
    if (value) vval[at] = value[j];
    ++at;
  }
}
// the keys that came with an embedding, in any order (the per-key passes below only visit these)
__global__ void dfc_keylist_kernel(size_t U, const unsigned char* __restrict__ has, unsigned* count, uint32_t* __restrict__ list) {
  const size_t u = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (u < U && has[u]) list[atomicAdd(count, 1u)] = (uint32_t)u;
}

// learn/difacto/loss.h:66-75: XV[i,:] = sum_j V[idx_j,:] val_j and xxvv[i,:] = sum_j V[idx_j,:]^2 val_j^2
// over the entries whose key carries an embedding, in CSR order. One thread per (row, d).
__global__ void __launch_bounds__(256)
dfc_xv_kernel(size_t rows, const unsigned* __restrict__ vstart, const unsigned* __restrict__ vcnt,
              const uint32_t* __restrict__ vidx, const float* __restrict__ vval, bool valued, const float* __restrict__ V,
              int dim, float* __restrict__ XV, float* __restrict__ xxvv) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t r = t / dim;
  const int d = (int)(t % dim);
  if (r >= rows) return;
  float a = 0.f, b = 0.f;
  const unsigned j0 = vstart[r], j1 = j0 + vcnt[r];
  for (unsigned j = j0; j < j1; ++j) {
    const float v = __ldg(V + (size_t)vidx[j] * dim + d);
    const float vv = v * v;
    if (valued) { const float x = vval[j]; a += v * x; b += vv * (x * x); }
    else { a += v; b += vv; }
  }
  XV[t] = a; xxvv[t] = b;
}

// learn/difacto/loss.h:76-83: py[i] += .5 * sum_d (XV^2 - xxvv)   (double product, float store)
__global__ void dfc_combine_kernel(size_t rows, int dim, const float* __restrict__ XV, const float* __restrict__ xxvv,
                                   float* py) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows) return;
  float s = 0.f;
  for (int d = 0; d < dim; ++d) { const float t = XV[i * dim + d]; s += t * t - xxvv[i * dim + d]; }
  py[i] = (float)((double)py[i] + .5 * (double)s);
}

// learn/difacto/loss.h:134-140: XV[i,:] *= p_i
__global__ void dfc_scale_rows_kernel(size_t rows, int dim, const float* __restrict__ p, float* XV) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < rows * dim) XV[t] *= p[t / dim];
}

// learn/difacto/loss.h:142 accumulation part: G[idx_j, d] += XVp[i, d] * val_j, and
// :121-123: xxp[idx_j] += p_i * val_j^2, for entries whose key carries an embedding. Fixed point.
__global__ void __launch_bounds__(256)
dfc_trans_kernel(size_t rows, const unsigned* __restrict__ vstart, const unsigned* __restrict__ vcnt,
                 const uint32_t* __restrict__ vidx, const float* __restrict__ vval, bool valued, int dim,
                 const float* __restrict__ XVp, const float* __restrict__ p, long long* accG, long long* accX,
                 const FixedScale* __restrict__ fsG, const FixedScale* __restrict__ fsX) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t r = t / dim;
  const int d = (int)(t % dim);
  if (r >= rows) return;
  const unsigned j0 = vstart[r], j1 = j0 + vcnt[r];
  if (j0 == j1) return;
  const float x = XVp[t], pi = p[r];
  const float sg = fsG->scale, sx = fsX->scale;
  for (unsigned j = j0; j < j1; ++j) {
    const uint32_t c = vidx[j];
    const float v = valued ? vval[j] : 1.f;
    fixed_add(accG + (size_t)c * dim + d, valued ? x * v : x, sg);
    if (d == 0) fixed_add(accX + c, valued ? pi * (v * v) : pi, sx);
  }
}

// Gradient dropout (learn/difacto/loss.h:150-154): `if ((T)rand() / RAND_MAX > 1 - V.dropout) g = 0` for every float of
// V.weight in order -- V.weight holds the keys that came with an embedding, ascending, dim floats each. The rand()
// values are the caller's (wh_difacto_set_dropout_stream): float number pos[u] * dim + d of this minibatch takes value
// number *cursor + pos[u] * dim + d of the stream, pos[u] = the key's rank among the keys with an embedding.
struct DfcDropout {
  float dropout = 0.f;
  const int* stream = nullptr; size_t len = 0;
  long long* cursor = nullptr;       // device: values consumed so far
  unsigned* pos = nullptr;           // device scratch, U entries
  unsigned* total = nullptr;         // device: keys with an embedding in this minibatch
};

// G[u,:] = (-xxp[u] * V[u,:]) + sum (learn/difacto/loss.h:126-131,142), then clip (:145-148), then dropout (:150-154).
// Also the sum of squares for the optional normalisation (:160-166), in fp64 per block.
__global__ void __launch_bounds__(256)
dfc_gv_finish_kernel(const unsigned* __restrict__ nkeys, const uint32_t* __restrict__ klist, int dim,
                     const float* __restrict__ V, long long* accG, long long* accX, const FixedScale* __restrict__ fsG,
                     const FixedScale* __restrict__ fsX, float clip, DfcDropout drop, float* __restrict__ G, double* sumsq) {
  const size_t n = (size_t)*nkeys * dim;
  double sq = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const size_t u = klist[i / dim];
    const int d = (int)(i % dim);
    const size_t t = u * dim + d;
    const long long ag = accG[t];
    accG[t] = 0;
    const float xxp = (float)((double)accX[u] * fsX->inv);
    const float z = V[t] * -xxp;
    float g = z * 1.0f + (float)((double)ag * fsG->inv);
    if (clip > 0.f) g = g > clip ? clip : (g < -clip ? -clip : g);
    if (drop.stream) {
      const unsigned long long o = (unsigned long long)*drop.cursor + (unsigned long long)drop.pos[u] * dim + d;
      if ((float)drop.stream[o % drop.len] / (float)2147483647 > 1 - drop.dropout) g = 0.f;
    }
    sq += (double)(g * g);
    G[t] = g;
  }
  __shared__ double s[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_down_sync(0xffffffffu, sq, o);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = sq;
  __syncthreads();
  if (threadIdx.x == 0) { double a = 0; for (int w = 0; w < 8; ++w) a += s[w]; if (a != 0) atomicAdd(sumsq, a); }
}
// after the pass above: the (X o X)^T p sums of the visited keys back to zero
__global__ void dfc_clear_keys_kernel(const unsigned* __restrict__ nkeys, const uint32_t* __restrict__ klist, long long* accX,
                                      double* sumsq) {
  const size_t n = *nkeys;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) accX[klist[i]] = 0;
  if (blockIdx.x == 0 && threadIdx.x == 0 && sumsq) *sumsq = 0;
}
// learn/difacto/loss.h:160-166
__global__ void dfc_normalize_kernel(const unsigned* __restrict__ nkeys, const uint32_t* __restrict__ klist, int dim, float* G,
                                     const double* sumsq) {
  const float nm = (float)*sumsq;
  if (nm < 1e-10f) return;
  const float r = sqrtf(nm);
  const size_t n = (size_t)*nkeys * dim;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const size_t t = (size_t)klist[i / dim] * dim + i % dim;
    G[t] = G[t] / r;
  }
}

struct HasVal { const unsigned char* has; __device__ unsigned operator()(size_t i) const { return has[i] ? 1u : 0u; } };
struct RankOut { unsigned* pos; __device__ void operator()(size_t i, unsigned, unsigned prefix) const { pos[i] = prefix; } };
__global__ void dfc_advance_kernel(long long* cursor, const unsigned* total, int dim) { *cursor += (long long)*total * dim; }

// Scratch of one DiFacto evaluation (all device pointers; sized by the caller).
struct DfcWork {
  float *py, *dual, *XV, *xxvv, *gw, *G;
  long long *accG, *accX;        // U*dim and U, zero between calls
  FixedScale *fsW, *fsG, *fsX;
  double* sumsq;
  EvalOut *evW, *ev;             // objective before / after the embedding term
  // the column subset for V (dfc_subset_kernel) and the keys that came with an embedding
  unsigned *vstart, *vcnt, *vcursor, *kcount;
  uint32_t *vidx, *klist;
  float* vval;
};

// Loss::Evaluate then (if train) Loss::CalcGrad on the strided model {w, has, V}.
int difacto_loss_launch(whContext* ctx, whContext* auc_ctx, cudaEvent_t ev_fork, cudaEvent_t ev_auc, size_t rows,
                        size_t nnz, const size_t* offset, const uint32_t* index, const float* value,
                        const float* label, size_t U, int dim, const float* w, const unsigned char* has,
                        const float* V, bool any_v, float clip, float norm, bool train, const DfcWork& k,
                        const DfcDropout* drop = nullptr) {
  cudaStream_t st = ctx->stream;
  WH_CUDA(cudaMemsetAsync(k.py, 0, rows * 4, st));
  WH_TRY(spmv_times_launch(ctx, rows, offset, index, value, w, k.py));                       // py = X w
  WH_TRY(loss_forward_launch(ctx, WH_LOSS_LOGIT, rows, label, k.py, nullptr, k.evW, 0));     // objv_w
  const unsigned grd = (unsigned)div_up(rows * (size_t)(dim > 0 ? dim : 1), 256);
  const unsigned kgrid = (unsigned)ctx->num_sms * 8;
  if (any_v) {
    WH_CUDA(cudaMemsetAsync(k.vcursor, 0, 8, st));       // vcursor, kcount
    WH_LAUNCH(dfc_subset_kernel, (unsigned)div_up(rows, 256), 256, 0, st, rows, offset, index, value, has, k.vcursor, k.vstart,
              k.vcnt, k.vidx, k.vval);
    WH_LAUNCH(dfc_xv_kernel, grd, 256, 0, st, rows, k.vstart, k.vcnt, k.vidx, k.vval, value != nullptr, V, dim, k.XV, k.xxvv);
    WH_LAUNCH(dfc_combine_kernel, (unsigned)div_up(rows, 256), 256, 0, st, rows, dim, k.XV, k.xxvv, k.py);
  }
  WH_TRY(loss_forward_launch(ctx, WH_LOSS_LOGIT, rows, label, k.py, train ? k.dual : nullptr, k.ev, 0));
  WH_CUDA(cudaEventRecord(ev_fork, st));
  WH_CUDA(cudaStreamWaitEvent(auc_ctx->stream, ev_fork, 0));
  WH_TRY(auc_launch(auc_ctx, rows, label, k.py, k.ev));
  WH_CUDA(cudaEventRecord(ev_auc, auc_ctx->stream));
  if (train) {
    long long* accW;
    WH_TRY(ensure_acc(ctx, U, &accW));
    WH_TRY(spmv_trans_accumulate_launch(ctx, rows, nnz, offset, index, value, k.dual, accW, k.fsW, acc_copies(), U));
    WH_TRY(fixed_finish_launch(ctx, accW, U, k.fsW, k.gw, 1.f, acc_copies(), U));             // grad_w = X' p
    if (any_v) {
      WH_LAUNCH(dfc_scale_rows_kernel, grd, 256, 0, st, rows, dim, k.dual, k.XV);
      // scales: max|XVp| * max|val| for G, max|p| * max|val|^2 for xxp
      WH_TRY(fixed_prepare(ctx, k.fsG, k.XV, rows * dim, value, nnz, 1));
      WH_TRY(fixed_prepare(ctx, k.fsX, k.dual, rows, value, nnz, 2));
      WH_LAUNCH(dfc_trans_kernel, grd, 256, 0, st, rows, k.vstart, k.vcnt, k.vidx, k.vval, value != nullptr, dim, k.XV, k.dual,
                k.accG, k.accX, k.fsG, k.fsX);
      WH_LAUNCH(dfc_keylist_kernel, (unsigned)div_up(U ? U : 1, 256), 256, 0, st, U, has, k.kcount, k.klist);
      DfcDropout dr;
      if (drop && drop->stream && drop->dropout > 0.f && U) {
        dr = *drop;
        WH_TRY(ctx->scan_tmp.reserve(div_up(U, kScanTile) * sizeof(unsigned) + 256));
        WH_TRY(scan_and_apply<unsigned>(ctx, HasVal{has}, RankOut{dr.pos}, U, ctx->scan_tmp.as<unsigned>(), dr.total));
      }
      WH_LAUNCH(dfc_gv_finish_kernel, kgrid, 256, 0, st, k.kcount, k.klist, dim, V, k.accG, k.accX, k.fsG, k.fsX, clip, dr, k.G,
                k.sumsq);
      if (dr.stream) WH_LAUNCH(dfc_advance_kernel, 1, 1, 0, st, dr.cursor, dr.total, dim);
      if (norm != 0.f) WH_LAUNCH(dfc_normalize_kernel, kgrid, 256, 0, st, k.kcount, k.klist, dim, k.G, k.sumsq);
      WH_LAUNCH(dfc_clear_keys_kernel, kgrid, 256, 0, st, k.kcount, k.klist, k.accX, k.sumsq);
    }
  }
  WH_CUDA(cudaStreamWaitEvent(st, ev_auc, 0));
  return WH_OK;
}

}  // namespace wh

// ============================================================================ the server ==
// AdaGradEntry / AdaGradHandle of learn/difacto/async_sgd.h:135-297 on a dense shard.

struct whDifacto {
  whContext* ctx = nullptr;
  whDifactoConfig c;
  float v_alpha = .01f, v_beta = 1.f;
  wh::DevBuf base, hasv, V, cgV, counters, init_stream, drop_stream, drop_pos;
  size_t init_len = 0;
  size_t drop_len = 0;               // the worker's rand() values for gradient dropout (0: none given)
  bool drop_shared = false;          // ... or: dropout draws from the init stream (worker and server in one process)
  // worker side (one minibatch at a time)
  whContext ax;                      // AUC side stream
  cudaEvent_t ev_fork = nullptr, ev_auc = nullptr;
  wh::DevBuf in_index, in_offset, in_value, in_label, uniq, cnt, index, w, has, Vs, XV, xxvv, py, dual, gw, G,
      accG, accX, flags, scal, vrows, vidx, vval, klist;
  wh::PinnedBuf pinned;
  long long seen_new_w = 0, seen_new_V = 0;
  // between wh_difacto_begin and wh_difacto_finish (sharded jobs: the model lives elsewhere)
  bool b_open = false;
  size_t b_rows = 0, b_nnz = 0, b_U = 0;
  const size_t* b_offset = nullptr; const float* b_value = nullptr; const float* b_label = nullptr;
  size_t entries() const { return (size_t)(c.key_end - c.key_begin); }
};

namespace wh {

struct DfcParams {
  float l1, l2, alpha, beta, v_l2, v_alpha, v_beta, v_min, v_max;
  int dim, l1_shrk;
  unsigned thr;
  uint64_t key_begin, key_end, seed;
};
// counters: [0] new_w, [1] new_V, [2] keys out of range, [3] cursor into the init stream

__device__ __forceinline__ bool dfc_slot(const DfcParams& p, uint64_t k, uint64_t* slot, long long* counters) {
  if (k < p.key_begin || k >= p.key_end) { atomicAdd(reinterpret_cast<unsigned long long*>(counters + 2), 1ull); return false; }
  *slot = k - p.key_begin;
  return true;
}
// AdaGradHandle::Resize condition, async_sgd.h:249-250
__device__ __forceinline__ bool dfc_wants_v(const DfcParams& p, unsigned fea_cnt, bool has, float w0) {
  return p.dim > 0 && fea_cnt > p.thr && !has && (!p.l1_shrk || w0 != 0.f);
}

// Push with push_count (async_sgd.h:217-219): fea_cnt += (unsigned) recv[0]; Resize(val)
__global__ void dfc_count_push_kernel(DfcParams p, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ cnt,
                                      const unsigned* d_n, float4* base, const unsigned char* hasv, unsigned char* flags,
                                      long long* counters) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_n) return;
  uint64_t s;
  flags[i] = 0;
  if (!dfc_slot(p, keys[i], &s, counters)) return;
  float4 e = base[s];
  const unsigned fc = __float_as_uint(e.w) + cnt[i];
  e.w = __uint_as_float(fc);
  base[s] = e;
  flags[i] = dfc_wants_v(p, fc, hasv[s] != 0, e.x) ? 1 : 0;
}

// ZVPull, AdaGradHandle::Pull (async_sgd.h:234-244), into the strided minibatch model
__global__ void dfc_pull_kernel(DfcParams p, const uint64_t* __restrict__ keys, const unsigned* d_n,
                                const float4* __restrict__ base, const unsigned char* __restrict__ hasv,
                                const float* __restrict__ V, float* w, unsigned char* has, float* Vs, long long* counters) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_n) return;
  uint64_t s;
  if (!dfc_slot(p, keys[i], &s, counters)) { w[i] = 0.f; has[i] = 0; return; }
  const float w0 = base[s].x;
  const bool full = hasv[s] && !(p.l1_shrk && w0 == 0.f);
  w[i] = w0;
  has[i] = full ? 1 : 0;
  if (full) for (int j = 0; j < p.dim; ++j) Vs[i * p.dim + j] = V[s * p.dim + j];
}

// ZVPush, Push without push_count (async_sgd.h:220-231): UpdateW (:262-286) then UpdateV (:289-296) for the
// keys that were pulled with their embedding. flags[i] = this key must grow an embedding now.
__global__ void dfc_grad_push_kernel(DfcParams p, const uint64_t* __restrict__ keys, const unsigned* d_n,
                                     const float* __restrict__ gw, const float* __restrict__ G,
                                     const unsigned char* __restrict__ has_pulled, float4* base, unsigned char* hasv,
                                     float* V, float* cgV, unsigned char* flags, long long* counters) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *d_n) return;
  uint64_t s;
  flags[i] = 0;
  if (!dfc_slot(p, keys[i], &s, counters)) return;
  float4 e = base[s];   // {w0, z0, cg0, fea_cnt}
  const float w = e.x;
  float g = gw[i];
  g += p.l2 * w;
  const float cg = e.z;
  const float cg_new = sqrtf(cg * cg + g * g);
  e.z = cg_new;
  e.y -= g - (cg_new - cg) / p.alpha * w;
  const float z = e.y;
  if (z <= p.l1 && z >= -p.l1) e.x = 0.f;
  else { const float eta = (p.beta + cg_new) / p.alpha; e.x = (z > 0.f ? z - p.l1 : z + p.l1) / eta; }
  base[s] = e;
  if (w == 0.f && e.x != 0.f) {
    atomicAdd(reinterpret_cast<unsigned long long*>(counters), 1ull);
    flags[i] = dfc_wants_v(p, __float_as_uint(e.w), hasv[s] != 0, e.x) ? 1 : 0;
  } else if (w != 0.f && e.x == 0.f) {
    atomicAdd(reinterpret_cast<unsigned long long*>(counters), (unsigned long long)(-1ll));
  }
  if (has_pulled[i]) {
    for (int j = 0; j < p.dim; ++j) {
      float* vw = V + s * p.dim + j;
      float* vc = cgV + s * p.dim + j;
      const float grad = G[i * p.dim + j] + p.v_l2 * *vw;
      const float c2 = sqrtf(*vc * *vc + grad * grad);
      *vc = c2;
      const float eta = p.v_alpha / (c2 + p.v_beta);
      *vw -= eta * grad;
    }
  }
}

// Embedding creation in message (= ascending key) order, async_sgd.h:251-257. With an init stream the o-th
// new float takes the o-th unread value of the caller's glibc rand() sequence (the reference's arrival-order
// semantics); without one, a counter-based hash of (seed, key, j).
struct GrowFlag {
  const unsigned char* flags; const unsigned* d_n;
  __device__ unsigned operator()(size_t i) const { return i < *d_n ? flags[i] : 0u; }
};
struct GrowApply {
  DfcParams p; const uint64_t* keys; unsigned char* hasv; float* V; float* cgV;
  const int* stream; size_t stream_len; long long* counters;
  __device__ void operator()(size_t i, unsigned v, unsigned ordinal) const {
    if (!v) return;
    const uint64_t s = keys[i] - p.key_begin;
    hasv[s] = 1;
    const unsigned long long cur = (unsigned long long)counters[3];
    for (int j = 0; j < p.dim; ++j) {
      int r;
      if (stream) r = stream[(cur + (unsigned long long)ordinal * p.dim + j) % stream_len];
      else {
        unsigned long long h = (keys[i] + 0x9E3779B97F4A7C15ull * (p.seed + 1)) ^ ((unsigned long long)j << 56);
        h ^= h >> 30; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 27; h *= 0x94D049BB133111EBull; h ^= h >> 31;
        r = (int)(h >> 33);
      }
      V[s * p.dim + j] = (float)r / (float)2147483647 * (p.v_max - p.v_min) + p.v_min;
      cgV[s * p.dim + j] = 0.f;
    }
    atomicAdd(reinterpret_cast<unsigned long long*>(counters + 1), (unsigned long long)p.dim);
  }
};
__global__ void dfc_advance_cursor_kernel(long long* counters, const unsigned* total, int dim) {
  counters[3] += (long long)*total * dim;
}

static DfcParams dfc_params(const whDifacto* d) {
  DfcParams p;
  p.l1 = d->c.lambda_l1; p.l2 = d->c.lambda_l2; p.alpha = d->c.lr_eta; p.beta = d->c.lr_beta;
  p.v_l2 = d->c.v_lambda_l2; p.v_alpha = d->v_alpha; p.v_beta = d->v_beta;
  p.v_min = -d->c.init_scale; p.v_max = d->c.init_scale;
  p.dim = d->c.dim; p.l1_shrk = d->c.l1_shrk; p.thr = (unsigned)d->c.threshold;
  p.key_begin = d->c.key_begin; p.key_end = d->c.key_end; p.seed = d->c.seed;
  return p;
}

static int dfc_grow(whDifacto* d, const uint64_t* keys, size_t ucap, const unsigned* d_n) {
  whContext* ctx = d->ctx;
  if (d->c.dim == 0) return WH_OK;
  WH_TRY(ctx->scan_tmp.reserve(div_up(ucap, kScanTile) * sizeof(unsigned) + 256));
  unsigned* total = d->scal.as<unsigned>() + 8;
  GrowFlag f{d->flags.as<unsigned char>(), d_n};
  GrowApply a{dfc_params(d), keys, d->hasv.as<unsigned char>(), d->V.as<float>(), d->cgV.as<float>(),
              d->init_len ? d->init_stream.as<int>() : nullptr, d->init_len, d->counters.as<long long>()};
  WH_TRY(scan_and_apply<unsigned>(ctx, f, a, ucap, ctx->scan_tmp.as<unsigned>(), total));
  WH_LAUNCH(dfc_advance_cursor_kernel, 1, 1, 0, ctx->stream, d->counters.as<long long>(), total, d->c.dim);
  return WH_OK;
}

}  // namespace wh

// ================================================================================ C ABI ==
using namespace wh;

namespace {

// model payload <-> strided form (learn/difacto/loss.h:185-218 Data::Load, :256-264 Data::Save)
__global__ void dfc_unpack_kernel(size_t U, int dim, const float* __restrict__ model, const int* __restrict__ siz,
                                  const unsigned* __restrict__ pos, float* w, unsigned char* has, float* Vs) {
  const size_t u = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  const int s = siz[u];
  w[u] = s > 0 ? model[pos[u]] : 0.f;
  has[u] = (dim > 0 && s == dim + 1) ? 1 : 0;
  if (has[u]) for (int j = 0; j < dim; ++j) Vs[u * dim + j] = model[pos[u] + 1 + j];
}
__global__ void dfc_pack_kernel(size_t U, int dim, const int* __restrict__ siz, const unsigned* __restrict__ pos,
                                const float* __restrict__ gw, const unsigned char* __restrict__ has,
                                const float* __restrict__ G, float* grad) {
  const size_t u = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  if (siz[u] > 0) grad[pos[u]] = gw[u];
  if (has[u]) for (int j = 0; j < dim; ++j) grad[pos[u] + 1 + j] = G[u * dim + j];
}
struct SizVal { const int* siz; __device__ unsigned operator()(size_t i) const { return (unsigned)siz[i]; } };
struct PosOut { unsigned* pos; __device__ void operator()(size_t i, unsigned, unsigned prefix) const { pos[i] = prefix; } };

__global__ void dfc_state_kernel(DfcParams p, const uint64_t* __restrict__ keys, size_t n, const float4* __restrict__ base,
                                 const unsigned char* __restrict__ hasv, const float* __restrict__ V,
                                 const float* __restrict__ cgV, uint32_t* fea_cnt, int* size, float* w, float* cg,
                                 long long* counters) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t s;
  const int dim = p.dim;
  for (int j = 0; j < dim + 1; ++j) w[i * (dim + 1) + j] = 0.f;
  for (int j = 0; j < dim + 2; ++j) cg[i * (dim + 2) + j] = 0.f;
  fea_cnt[i] = 0; size[i] = 1;
  if (!dfc_slot(p, keys[i], &s, counters)) return;
  const float4 e = base[s];
  fea_cnt[i] = __float_as_uint(e.w);
  w[i * (dim + 1)] = e.x; cg[i * (dim + 2)] = e.z; cg[i * (dim + 2) + 1] = e.y;
  if (hasv[s]) {
    size[i] = dim + 1;
    for (int j = 0; j < dim; ++j) { w[i * (dim + 1) + 1 + j] = V[s * dim + j]; cg[i * (dim + 2) + 2 + j] = cgV[s * dim + j]; }
  }
}

int dfc_work(whDifacto* d, size_t rows, size_t U, size_t nnz, DfcWork* k) {
  const int dim = d->c.dim > 0 ? d->c.dim : 1;
  if (d->c.dim > 0) {
    WH_TRY(d->vrows.reserve(rows * 8)); WH_TRY(d->vidx.reserve((nnz ? nnz : 1) * 4)); WH_TRY(d->vval.reserve((nnz ? nnz : 1) * 4));
    WH_TRY(d->klist.reserve((U ? U : 1) * 4));
  }
  k->vstart = d->vrows.as<unsigned>(); k->vcnt = k->vstart + rows;
  k->vidx = d->vidx.as<uint32_t>(); k->vval = d->vval.as<float>(); k->klist = d->klist.as<uint32_t>();
  WH_TRY(d->py.reserve(rows * 4)); WH_TRY(d->dual.reserve(rows * 4));
  WH_TRY(d->XV.reserve(rows * dim * 4)); WH_TRY(d->xxvv.reserve(rows * dim * 4));
  WH_TRY(d->gw.reserve((U ? U : 1) * 4)); WH_TRY(d->G.reserve((U ? U : 1) * dim * 4));
  if (d->accG.cap < (U ? U : 1) * dim * 8) {
    WH_TRY(d->accG.reserve((U ? U : 1) * dim * 8));
    WH_CUDA(cudaMemsetAsync(d->accG.p, 0, d->accG.cap, d->ctx->stream));
  }
  if (d->accX.cap < (U ? U : 1) * 8) {
    WH_TRY(d->accX.reserve((U ? U : 1) * 8));
    WH_CUDA(cudaMemsetAsync(d->accX.p, 0, d->accX.cap, d->ctx->stream));
  }
  char* sc = d->scal.as<char>();
  k->py = d->py.as<float>(); k->dual = d->dual.as<float>(); k->XV = d->XV.as<float>(); k->xxvv = d->xxvv.as<float>();
  k->gw = d->gw.as<float>(); k->G = d->G.as<float>();
  k->accG = d->accG.as<long long>(); k->accX = d->accX.as<long long>();
  k->fsW = reinterpret_cast<FixedScale*>(sc + 256); k->fsG = reinterpret_cast<FixedScale*>(sc + 320);
  k->fsX = reinterpret_cast<FixedScale*>(sc + 384); k->sumsq = reinterpret_cast<double*>(sc + 448);
  k->evW = reinterpret_cast<EvalOut*>(sc + 512); k->ev = reinterpret_cast<EvalOut*>(sc + 768);
  k->vcursor = reinterpret_cast<unsigned*>(sc + 64); k->kcount = k->vcursor + 1;
  return WH_OK;
}

void dfc_prog(const EvalOut* evW, const EvalOut* ev, size_t rows, double* prog) {
  EvalHost a = eval_finish(*evW, rows), b = eval_finish(*ev, rows);
  prog[0] = b.objv; prog[1] = b.auc; prog[2] = a.objv; prog[3] = 0; prog[4] = 1; prog[5] = (double)rows;
  prog[6] = 0; prog[7] = 0;
}

}  // namespace

extern "C" {

int wh_difacto_create(whContext* ctx, const whDifactoConfig* c, whDifacto** out) {
  WH_REQUIRE(ctx && c && out, "wh_difacto_create: NULL argument");
  WH_REQUIRE(c->dim >= 0 && c->dim <= 256, "wh_difacto_create: dim %d outside [0, 256]", c->dim);
  WH_REQUIRE(c->dropout >= 0.f && c->dropout <= 1.f, "wh_difacto_create: dropout %g outside [0, 1]", (double)c->dropout);
  WH_REQUIRE(c->key_end > c->key_begin && c->key_end - c->key_begin <= (1ull << 30), "wh_difacto_create: bad key range");
  WH_REQUIRE(c->max_key != 0 && c->lr_eta > 0, "wh_difacto_create: max_key and lr_eta must be positive");
  DeviceGuard g(ctx->device);
  whDifacto* d = new whDifacto();
  d->ctx = ctx; d->c = *c;
  d->v_alpha = c->v_has_lr_eta ? c->v_lr_eta : c->lr_eta;     // async_sgd.h:322-323
  d->v_beta = c->v_has_lr_beta ? c->v_lr_beta : c->lr_beta;
  const size_t n = d->entries(), dim = c->dim > 0 ? c->dim : 1;
  int r = d->base.reserve(n * 16);
  if (r == WH_OK) r = d->hasv.reserve(n);
  if (r == WH_OK && c->dim > 0) r = d->V.reserve(n * dim * 4);
  if (r == WH_OK && c->dim > 0) r = d->cgV.reserve(n * dim * 4);
  if (r == WH_OK) r = d->counters.reserve(64);
  if (r == WH_OK) r = d->scal.reserve(4096);
  if (r == WH_OK) r = d->pinned.reserve(4096);
  if (r != WH_OK) { delete d; return r; }
  WH_CUDA(cudaMemsetAsync(d->scal.p, 0, 4096, ctx->stream));
  WH_CUDA(cudaMemsetAsync(d->base.p, 0, d->base.cap, ctx->stream));
  WH_CUDA(cudaMemsetAsync(d->hasv.p, 0, d->hasv.cap, ctx->stream));
  WH_CUDA(cudaMemsetAsync(d->counters.p, 0, 64, ctx->stream));
  d->ax.device = ctx->device; d->ax.num_sms = ctx->num_sms;
  WH_CUDA(cudaStreamCreateWithFlags(&d->ax.stream, cudaStreamNonBlocking));
  WH_TRY(d->ax.small.reserve(1 << 14));
  WH_CUDA(cudaEventCreateWithFlags(&d->ev_fork, cudaEventDisableTiming));
  WH_CUDA(cudaEventCreateWithFlags(&d->ev_auc, cudaEventDisableTiming));
  *out = d;
  return WH_OK;
}

int wh_difacto_destroy(whDifacto* d) {
  if (!d) return WH_OK;
  DeviceGuard g(d->ctx->device);
  cudaStreamSynchronize(d->ctx->stream);
  cudaStreamSynchronize(d->ax.stream);
  for (wh::DevBuf* b : {&d->base, &d->hasv, &d->V, &d->cgV, &d->counters, &d->init_stream, &d->drop_stream, &d->drop_pos, &d->in_index, &d->in_offset,
                        &d->in_value, &d->in_label, &d->uniq, &d->cnt, &d->index, &d->w, &d->has, &d->Vs, &d->XV,
                        &d->xxvv, &d->py, &d->dual, &d->gw, &d->G, &d->accG, &d->accX, &d->flags, &d->scal, &d->vrows, &d->vidx, &d->vval, &d->klist,
                        &d->ax.small, &d->ax.scan_tmp})
    b->release();
  for (auto& b : d->ax.sort_tmp) b.release();
  for (auto& b : d->ax.stage) b.release();
  d->pinned.release();
  cudaEventDestroy(d->ev_fork); cudaEventDestroy(d->ev_auc);
  cudaStreamDestroy(d->ax.stream);
  delete d;
  return WH_OK;
}

int wh_difacto_set_init_stream(whDifacto* d, const int* v, size_t n) {
  WH_REQUIRE(d && (n == 0 || v), "wh_difacto_set_init_stream: NULL argument");
  DeviceGuard g(d->ctx->device);
  WH_CUDA(cudaStreamSynchronize(d->ctx->stream));
  d->init_len = n;
  if (n) {
    WH_TRY(d->init_stream.reserve(n * 4));
    WH_CUDA(cudaMemcpyAsync(d->init_stream.p, v, n * 4, cudaMemcpyHostToDevice, d->ctx->stream));
  }
  WH_CUDA(cudaMemsetAsync(d->counters.as<long long>() + 3, 0, 8, d->ctx->stream));
  WH_CUDA(cudaStreamSynchronize(d->ctx->stream));
  return WH_OK;
}

int wh_difacto_set_dropout_stream(whDifacto* d, const int* v, size_t n, int shared_with_init) {
  WH_REQUIRE(d && (shared_with_init || n == 0 || v), "wh_difacto_set_dropout_stream: NULL argument");
  DeviceGuard g(d->ctx->device);
  WH_CUDA(cudaStreamSynchronize(d->ctx->stream));
  d->drop_shared = shared_with_init != 0;
  d->drop_len = d->drop_shared ? 0 : n;
  if (d->drop_len) {
    WH_TRY(d->drop_stream.reserve(n * 4));
    WH_CUDA(cudaMemcpyAsync(d->drop_stream.p, v, n * 4, cudaMemcpyHostToDevice, d->ctx->stream));
  }
  WH_CUDA(cudaMemsetAsync(d->counters.as<long long>() + 4, 0, 8, d->ctx->stream));
  WH_CUDA(cudaStreamSynchronize(d->ctx->stream));
  return WH_OK;
}

// the dropout arguments of one training minibatch with U keys (none: conf.dropout == 0 or no embeddings)
static int dfc_dropout(whDifacto* d, size_t U, bool train, DfcDropout* out) {
  *out = DfcDropout();
  if (!train || d->c.dropout <= 0.f || d->c.dim == 0) return WH_OK;
  const bool shared = d->drop_shared && d->init_len;
  WH_REQUIRE(shared || d->drop_len, "wh_difacto: embedding dropout = %g draws from the worker's rand() (learn/difacto/loss.h:"
             "150-154): hand over that sequence with wh_difacto_set_dropout_stream before training", (double)d->c.dropout);
  WH_TRY(d->drop_pos.reserve((U ? U : 1) * 4));
  out->dropout = d->c.dropout;
  out->stream = shared ? d->init_stream.as<int>() : d->drop_stream.as<int>();
  out->len = shared ? d->init_len : d->drop_len;
  out->cursor = d->counters.as<long long>() + (shared ? 3 : 4);
  out->pos = d->drop_pos.as<unsigned>();
  out->total = d->scal.as<unsigned>() + 10;
  return WH_OK;
}

int wh_difacto_process_minibatch(whDifacto* d, const whRowBlock* blk, int mem, int train, int data_pass, double* prog) {
  WH_REQUIRE(d && blk && prog, "wh_difacto_process_minibatch: NULL argument");
  WH_REQUIRE(mem == WH_HOST || mem == WH_DEVICE, "wh_difacto_process_minibatch: bad mem %d", mem);
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  for (int i = 0; i < 8; ++i) prog[i] = 0;
  const size_t rows = blk->size;
  if (rows == 0) return WH_OK;
  WH_REQUIRE(blk->offset && blk->index && blk->label, "wh_difacto_process_minibatch: offset/index/label is NULL");
  size_t n;
  if (mem == WH_HOST) n = blk->offset[rows];
  else {
    WH_CUDA(cudaMemcpyAsync(d->pinned.p, blk->offset + rows, sizeof(size_t), cudaMemcpyDeviceToHost, ctx->stream));
    WH_CUDA(cudaStreamSynchronize(ctx->stream));
    n = d->pinned.as<size_t>()[0];
  }
  WH_REQUIRE(n > 0 && n < (size_t)UINT32_MAX, "wh_difacto_process_minibatch: %zu entries", n);
  const int dim = d->c.dim, sd = dim > 0 ? dim : 1;
  const uint64_t* d_index; const size_t* d_offset; const float* d_value; const float* d_label;
  WH_TRY(to_device(ctx, d->in_index, blk->index, n, mem, &d_index));
  WH_TRY(to_device(ctx, d->in_offset, blk->offset, rows + 1, mem, &d_offset));
  WH_TRY(to_device(ctx, d->in_value, blk->value, n, mem, &d_value));
  WH_TRY(to_device(ctx, d->in_label, blk->label, rows, mem, &d_label));
  const size_t kcap = (size_t)std::min<uint64_t>(n, d->c.max_key);
  WH_TRY(d->uniq.reserve(kcap * 8)); WH_TRY(d->cnt.reserve(kcap * 4)); WH_TRY(d->index.reserve(n * 4));
  unsigned* d_nuniq = d->scal.as<unsigned>();
  // Localizer<FeaID>::Localize(mb, data, feaid, feacnt): async_sgd.h:368-371
  LocInput in{d_index, n, make_xform(d->c.max_key)};
  // (the counts are only pushed in the first pass, async_sgd.h:374: every later minibatch saves one atomic per entry)
  uint32_t* want_cnt = (train && data_pass == 0 && d->c.dim > 0) ? d->cnt.as<uint32_t>() : nullptr;
  if (direct_supported(d->c.max_key)) WH_TRY(localize_direct(ctx, in, d->uniq.as<uint64_t>(), want_cnt, d->index.as<uint32_t>(), d_nuniq));
  else WH_TRY(localize_sort(ctx, in, d->uniq.as<uint64_t>(), want_cnt, d->index.as<uint32_t>(), d_nuniq));
  // the per-key buffers of a factorization machine are dim times wider than the linear model's: size them
  // by the actual number of unique keys (one small read-back) rather than by its upper bound
  WH_CUDA(cudaMemcpyAsync(d->pinned.p, d_nuniq, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  const size_t ucap = d->pinned.as<unsigned>()[0];
  WH_TRY(d->w.reserve((ucap ? ucap : 1) * 4)); WH_TRY(d->has.reserve(ucap ? ucap : 1));
  WH_TRY(d->Vs.reserve((ucap ? ucap : 1) * sd * 4)); WH_TRY(d->flags.reserve(ucap ? ucap : 1));
  DfcWork k;
  WH_TRY(dfc_work(d, rows, ucap, n, &k));
  const DfcParams p = dfc_params(d);
  long long* counters = d->counters.as<long long>();
  const unsigned ugrid = (unsigned)div_up(ucap, 256);
  // the feature-count push of the first pass: async_sgd.h:374-381
  if (train && data_pass == 0 && dim > 0) {
    WH_LAUNCH(dfc_count_push_kernel, ugrid, 256, 0, ctx->stream, p, d->uniq.as<uint64_t>(), d->cnt.as<uint32_t>(), d_nuniq,
              d->base.as<float4>(), d->hasv.as<unsigned char>(), d->flags.as<unsigned char>(), counters);
    WH_TRY(dfc_grow(d, d->uniq.as<uint64_t>(), ucap, d_nuniq));
  }
  // ZVPull
  WH_LAUNCH(dfc_pull_kernel, ugrid, 256, 0, ctx->stream, p, d->uniq.as<uint64_t>(), d_nuniq, d->base.as<float4>(),
            d->hasv.as<unsigned char>(), d->V.as<float>(), d->w.as<float>(), d->has.as<unsigned char>(), d->Vs.as<float>(), counters);
  // Loss::Evaluate / CalcGrad. U is only known on the device: the per-key kernels run over ucap entries; the
  // tail beyond U is never referenced by the CSR, but must read as "no embedding, zero accumulators".
  DfcDropout drop;
  WH_TRY(dfc_dropout(d, ucap, train != 0, &drop));
  WH_TRY(difacto_loss_launch(ctx, &d->ax, d->ev_fork, d->ev_auc, rows, n, d_offset, d->index.as<uint32_t>(), d_value, d_label,
                             ucap, sd, d->w.as<float>(), d->has.as<unsigned char>(), d->Vs.as<float>(), dim > 0,
                             d->c.grad_clipping, d->c.grad_normalization, train != 0, k, &drop));
  if (train) {
    // ZVPush
    WH_LAUNCH(dfc_grad_push_kernel, ugrid, 256, 0, ctx->stream, p, d->uniq.as<uint64_t>(), d_nuniq, k.gw, k.G,
              d->has.as<unsigned char>(), d->base.as<float4>(), d->hasv.as<unsigned char>(), d->V.as<float>(),
              d->cgV.as<float>(), d->flags.as<unsigned char>(), counters);
    WH_TRY(dfc_grow(d, d->uniq.as<uint64_t>(), ucap, d_nuniq));
  }
  char* pin = d->pinned.as<char>();
  WH_CUDA(cudaMemcpyAsync(pin, k.evW, sizeof(EvalOut), cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaMemcpyAsync(pin + 256, k.ev, sizeof(EvalOut), cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaMemcpyAsync(pin + 512, counters, 32, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  const long long* cn = reinterpret_cast<const long long*>(pin + 512);
  if (cn[2]) {
    WH_CUDA(cudaMemsetAsync(counters + 2, 0, 8, ctx->stream));
    set_error("wh_difacto_process_minibatch: %lld keys outside the shard's range", cn[2]);
    return WH_ERR_INVALID;
  }
  dfc_prog(reinterpret_cast<const EvalOut*>(pin), reinterpret_cast<const EvalOut*>(pin + 256), rows, prog);
  prog[6] = (double)(cn[0] - d->seen_new_w); prog[7] = (double)(cn[1] - d->seen_new_V);
  d->seen_new_w = cn[0]; d->seen_new_V = cn[1];
  return WH_OK;
}

// ---- the same minibatch in phases, for key-range shards on several GPUs -------------------------------------
// AsyncWorker::ProcessMinibatch (learn/difacto/async_sgd.h:363-425) cut at its three exchanges: the count push
// (:374-381), ZVPull (:424) and ZVPush (:408-411). begin / finish are the worker's halves; the shard_* calls are what
// the server does with one message (:216-244). The variable-length payloads of ZVPull / ZVPush travel as fixed-width
// rows {w, has, V[dim]} per key -- `has` is the length flag (model_siz = 1 or dim + 1).

static int dfc_count_to_device(whDifacto* d, size_t n, unsigned** d_n) {
  unsigned* slot = d->scal.as<unsigned>() + 12;
  const unsigned v = (unsigned)n;
  WH_CUDA(cudaMemcpyAsync(slot, &v, 4, cudaMemcpyHostToDevice, d->ctx->stream));   // (pageable source: staged before the call returns)
  *d_n = slot;
  return WH_OK;
}

int wh_difacto_begin(whDifacto* d, const whRowBlock* blk, int mem, const uint64_t** d_uniq, const uint32_t** d_cnt, size_t* num_uniq) {
  WH_REQUIRE(d && blk && d_uniq && d_cnt && num_uniq, "wh_difacto_begin: NULL argument");
  WH_REQUIRE(mem == WH_HOST || mem == WH_DEVICE, "wh_difacto_begin: bad mem %d", mem);
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  const size_t rows = blk->size;
  WH_REQUIRE(rows > 0 && blk->offset && blk->index && blk->label, "wh_difacto_begin: empty block or NULL offset/index/label");
  size_t n;
  if (mem == WH_HOST) n = blk->offset[rows];
  else {
    WH_CUDA(cudaMemcpyAsync(d->pinned.p, blk->offset + rows, sizeof(size_t), cudaMemcpyDeviceToHost, ctx->stream));
    WH_CUDA(cudaStreamSynchronize(ctx->stream));
    n = d->pinned.as<size_t>()[0];
  }
  WH_REQUIRE(n > 0 && n < (size_t)UINT32_MAX, "wh_difacto_begin: %zu entries", n);
  const uint64_t* d_index;
  WH_TRY(to_device(ctx, d->in_index, blk->index, n, mem, &d_index));
  WH_TRY(to_device(ctx, d->in_offset, blk->offset, rows + 1, mem, &d->b_offset));
  WH_TRY(to_device(ctx, d->in_value, blk->value, n, mem, &d->b_value));
  WH_TRY(to_device(ctx, d->in_label, blk->label, rows, mem, &d->b_label));
  const size_t kcap = (size_t)std::min<uint64_t>(n, d->c.max_key);
  WH_TRY(d->uniq.reserve(kcap * 8)); WH_TRY(d->cnt.reserve(kcap * 4)); WH_TRY(d->index.reserve(n * 4));
  unsigned* d_nuniq = d->scal.as<unsigned>();
  LocInput in{d_index, n, make_xform(d->c.max_key)};
  if (direct_supported(d->c.max_key)) WH_TRY(localize_direct(ctx, in, d->uniq.as<uint64_t>(), d->cnt.as<uint32_t>(), d->index.as<uint32_t>(), d_nuniq));
  else WH_TRY(localize_sort(ctx, in, d->uniq.as<uint64_t>(), d->cnt.as<uint32_t>(), d->index.as<uint32_t>(), d_nuniq));
  WH_CUDA(cudaMemcpyAsync(d->pinned.p, d_nuniq, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  d->b_rows = rows; d->b_nnz = n; d->b_U = d->pinned.as<unsigned>()[0];
  d->b_open = true;
  *d_uniq = d->uniq.as<uint64_t>(); *d_cnt = d->cnt.as<uint32_t>(); *num_uniq = d->b_U;
  return WH_OK;
}

int wh_difacto_finish(whDifacto* d, const float* d_w, const unsigned char* d_has, const float* d_V, int train, float* d_gw,
                      float* d_G, double* prog) {
  WH_REQUIRE(d && prog && d->b_open, "wh_difacto_finish: no minibatch begun");
  WH_REQUIRE(d->b_U == 0 || (d_w && d_has && (d->c.dim == 0 || d_V)), "wh_difacto_finish: the pulled model is NULL");
  WH_REQUIRE(!train || d->b_U == 0 || (d_gw && (d->c.dim == 0 || d_G)), "wh_difacto_finish: training needs gradient buffers");
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  d->b_open = false;
  const int dim = d->c.dim, sd = dim > 0 ? dim : 1;
  const size_t rows = d->b_rows, U = d->b_U;
  DfcWork k;
  WH_TRY(dfc_work(d, rows, U, d->b_nnz, &k));
  DfcDropout drop;
  WH_TRY(dfc_dropout(d, U, train != 0, &drop));
  WH_TRY(difacto_loss_launch(ctx, &d->ax, d->ev_fork, d->ev_auc, rows, d->b_nnz, d->b_offset, d->index.as<uint32_t>(), d->b_value,
                             d->b_label, U, sd, d_w, d_has, d_V, dim > 0, d->c.grad_clipping, d->c.grad_normalization, train != 0, k,
                             &drop));
  if (train && U) {
    WH_CUDA(cudaMemcpyAsync(d_gw, k.gw, U * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    if (dim > 0) WH_CUDA(cudaMemcpyAsync(d_G, k.G, U * (size_t)dim * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  }
  char* pin = d->pinned.as<char>();
  WH_CUDA(cudaMemcpyAsync(pin, k.evW, sizeof(EvalOut), cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaMemcpyAsync(pin + 256, k.ev, sizeof(EvalOut), cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  dfc_prog(reinterpret_cast<const EvalOut*>(pin), reinterpret_cast<const EvalOut*>(pin + 256), rows, prog);
  return WH_OK;
}

int wh_difacto_shard_count_push(whDifacto* d, const uint64_t* d_keys, const uint32_t* d_cnt, size_t n) {
  WH_REQUIRE(d && (n == 0 || (d_keys && d_cnt)), "wh_difacto_shard_count_push: NULL argument");
  if (n == 0 || d->c.dim == 0) return WH_OK;
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  WH_TRY(d->flags.reserve(n));
  unsigned* d_n;
  WH_TRY(dfc_count_to_device(d, n, &d_n));
  WH_LAUNCH(dfc_count_push_kernel, (unsigned)div_up(n, 256), 256, 0, ctx->stream, dfc_params(d), d_keys, d_cnt, d_n, d->base.as<float4>(),
            d->hasv.as<unsigned char>(), d->flags.as<unsigned char>(), d->counters.as<long long>());
  return dfc_grow(d, d_keys, n, d_n);
}

int wh_difacto_shard_pull(whDifacto* d, const uint64_t* d_keys, size_t n, float* d_w, unsigned char* d_has, float* d_V) {
  WH_REQUIRE(d && (n == 0 || (d_keys && d_w && d_has && (d->c.dim == 0 || d_V))), "wh_difacto_shard_pull: NULL argument");
  if (n == 0) return WH_OK;
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  unsigned* d_n;
  WH_TRY(dfc_count_to_device(d, n, &d_n));
  WH_LAUNCH(dfc_pull_kernel, (unsigned)div_up(n, 256), 256, 0, ctx->stream, dfc_params(d), d_keys, d_n, d->base.as<float4>(),
            d->hasv.as<unsigned char>(), d->V.as<float>(), d_w, d_has, d_V, d->counters.as<long long>());
  return WH_OK;
}

int wh_difacto_shard_push(whDifacto* d, const uint64_t* d_keys, size_t n, const float* d_gw, const float* d_G, const unsigned char* d_has) {
  WH_REQUIRE(d && (n == 0 || (d_keys && d_gw && d_has && (d->c.dim == 0 || d_G))), "wh_difacto_shard_push: NULL argument");
  if (n == 0) return WH_OK;
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  WH_TRY(d->flags.reserve(n));
  unsigned* d_n;
  WH_TRY(dfc_count_to_device(d, n, &d_n));
  WH_LAUNCH(dfc_grad_push_kernel, (unsigned)div_up(n, 256), 256, 0, ctx->stream, dfc_params(d), d_keys, d_n, d_gw, d_G, d_has,
            d->base.as<float4>(), d->hasv.as<unsigned char>(), d->V.as<float>(), d->cgV.as<float>(), d->flags.as<unsigned char>(),
            d->counters.as<long long>());
  return dfc_grow(d, d_keys, n, d_n);
}

// new_w / new_V of this shard since the last call (the progress numbers the servers own); also reports bad keys
int wh_difacto_shard_progress(whDifacto* d, double* new_w, double* new_V) {
  WH_REQUIRE(d && new_w && new_V, "wh_difacto_shard_progress: NULL argument");
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  char* pin = d->pinned.as<char>();
  WH_CUDA(cudaMemcpyAsync(pin + 512, d->counters.p, 32, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  const long long* cn = reinterpret_cast<const long long*>(pin + 512);
  if (cn[2]) {
    WH_CUDA(cudaMemsetAsync(d->counters.as<long long>() + 2, 0, 8, ctx->stream));
    set_error("wh_difacto: %lld keys outside the shard's range", cn[2]);
    return WH_ERR_INVALID;
  }
  *new_w = (double)(cn[0] - d->seen_new_w); *new_V = (double)(cn[1] - d->seen_new_V);
  d->seen_new_w = cn[0]; d->seen_new_V = cn[1];
  return WH_OK;
}

// ---- ps::KVWorker<float>::ZPush(cmd = kPushFeaCnt) / ZVPull / ZVPush on HOST arrays (the C++ facade, host/ps.h) ----
// One message each against this shard: keys ascending, payloads concatenated with a length per key, as ps-lite moves
// them (call sites learn/difacto/async_sgd.h:374-381,408-411,424). Synchronous.

// keys outside the shard's range seen by the kernels so far (stream already synchronised)
static int dfc_bad_keys(whDifacto* d) {
  long long bad = 0;
  WH_CUDA(cudaMemcpy(&bad, d->counters.as<long long>() + 2, 8, cudaMemcpyDeviceToHost));
  if (bad) {
    WH_CUDA(cudaMemset(d->counters.as<long long>() + 2, 0, 8));
    set_error("wh_difacto: %lld keys outside the shard's range", bad);
    return WH_ERR_INVALID;
  }
  return WH_OK;
}

static int dfc_host_keys(whDifacto* d, const uint64_t* keys, size_t n, const uint64_t** dk) {
  for (size_t i = 1; i < n; ++i)
    if (keys[i] <= keys[i - 1]) { set_error("wh_difacto: the keys of a message must ascend (position %zu)", i); return WH_ERR_INVALID; }
  return to_device(d->ctx, d->ctx->stage[0], keys, n, WH_HOST, dk);
}

int wh_difacto_zpush_count(whDifacto* d, const uint64_t* keys, const float* counts, size_t n) {
  WH_REQUIRE(d && (n == 0 || (keys && counts)), "wh_difacto_zpush_count: NULL argument");
  if (n == 0) return WH_OK;
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  const uint64_t* dk;
  WH_TRY(dfc_host_keys(d, keys, n, &dk));
  std::vector<uint32_t> c(n);
  for (size_t i = 0; i < n; ++i) c[i] = (unsigned)counts[i];      // val.fea_cnt += (unsigned) recv[0], async_sgd.h:218
  const uint32_t* dc;
  WH_TRY(to_device(ctx, ctx->stage[1], c.data(), n, WH_HOST, &dc));
  WH_TRY(wh_difacto_shard_count_push(d, dk, dc, n));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  return dfc_bad_keys(d);
}

int wh_difacto_zvpull(whDifacto* d, const uint64_t* keys, size_t n, float* vals, int* sizes, size_t* vals_len) {
  WH_REQUIRE(d && vals_len && (n == 0 || (keys && vals && sizes)), "wh_difacto_zvpull: NULL argument");
  *vals_len = 0;
  if (n == 0) return WH_OK;
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  const int dim = d->c.dim, sd = dim > 0 ? dim : 1;
  const uint64_t* dk;
  WH_TRY(dfc_host_keys(d, keys, n, &dk));
  WH_TRY(ctx->stage[1].reserve(n * 4)); WH_TRY(ctx->stage[2].reserve(n)); WH_TRY(ctx->stage[3].reserve(n * sd * 4));
  WH_TRY(wh_difacto_shard_pull(d, dk, n, ctx->stage[1].as<float>(), ctx->stage[2].as<unsigned char>(), ctx->stage[3].as<float>()));
  std::vector<float> w(n), V(dim > 0 ? n * (size_t)dim : 0);
  std::vector<unsigned char> has(n);
  WH_CUDA(cudaMemcpyAsync(w.data(), ctx->stage[1].p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaMemcpyAsync(has.data(), ctx->stage[2].p, n, cudaMemcpyDeviceToHost, ctx->stream));
  if (dim > 0) WH_CUDA(cudaMemcpyAsync(V.data(), ctx->stage[3].p, n * (size_t)dim * 4, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  WH_TRY(dfc_bad_keys(d));
  size_t o = 0;
  for (size_t i = 0; i < n; ++i) {       // AdaGradHandle::Pull, async_sgd.h:234-244: {w0} or {w0, V[dim]}
    vals[o++] = w[i];
    if (has[i]) { memcpy(vals + o, V.data() + i * (size_t)dim, (size_t)dim * 4); o += (size_t)dim; }
    sizes[i] = has[i] ? dim + 1 : 1;
  }
  *vals_len = o;
  return WH_OK;
}

int wh_difacto_zvpush(whDifacto* d, const uint64_t* keys, size_t n, const float* vals, const int* sizes, size_t vals_len) {
  WH_REQUIRE(d && (n == 0 || (keys && vals && sizes)), "wh_difacto_zvpush: NULL argument");
  if (n == 0) return WH_OK;
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  const int dim = d->c.dim;
  std::vector<float> gw(n), G(dim > 0 ? n * (size_t)dim : 0, 0.f);
  std::vector<unsigned char> has(n);
  size_t o = 0;
  for (size_t i = 0; i < n; ++i) {
    const int s = sizes[i];
    WH_REQUIRE(s == 1 || (dim > 0 && s == dim + 1), "wh_difacto_zvpush: key %zu carries %d floats (1 or dim + 1 = %d)", i, s, dim + 1);
    WH_REQUIRE(o + (size_t)s <= vals_len, "wh_difacto_zvpush: the lengths add up to more than the %zu values given", vals_len);
    gw[i] = vals[o];
    has[i] = s > 1;
    if (s > 1) memcpy(G.data() + i * (size_t)dim, vals + o + 1, (size_t)dim * 4);
    o += (size_t)s;
  }
  WH_REQUIRE(o == vals_len, "wh_difacto_zvpush: the lengths add up to %zu, %zu values given", o, vals_len);
  const uint64_t* dk; const float* dgw; const float* dG = nullptr; const unsigned char* dhas;
  WH_TRY(dfc_host_keys(d, keys, n, &dk));
  WH_TRY(to_device(ctx, ctx->stage[1], gw.data(), n, WH_HOST, &dgw));
  WH_TRY(to_device(ctx, ctx->stage[2], has.data(), n, WH_HOST, &dhas));
  if (dim > 0) WH_TRY(to_device(ctx, ctx->stage[3], G.data(), n * (size_t)dim, WH_HOST, &dG));
  WH_TRY(wh_difacto_shard_push(d, dk, n, dgw, dG, dhas));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  return dfc_bad_keys(d);
}

int wh_difacto_get_state(whDifacto* d, const uint64_t* keys, size_t n, uint32_t* fea_cnt, int* size, float* w, float* cg) {
  WH_REQUIRE(d && (n == 0 || (keys && fea_cnt && size && w && cg)), "wh_difacto_get_state: NULL argument");
  whContext* ctx = d->ctx;
  DeviceGuard g(ctx->device);
  if (n == 0) return WH_OK;
  const int dim = d->c.dim;
  const uint64_t* dk;
  WH_TRY(to_device(ctx, ctx->stage[0], keys, n, WH_HOST, &dk));
  WH_TRY(ctx->stage[1].reserve(n * 4)); WH_TRY(ctx->stage[2].reserve(n * 4));
  WH_TRY(ctx->stage[3].reserve(n * (dim + 1) * 4)); WH_TRY(ctx->stage[4].reserve(n * (dim + 2) * 4));
  WH_LAUNCH(dfc_state_kernel, (unsigned)div_up(n, 256), 256, 0, ctx->stream, dfc_params(d), dk, n, d->base.as<float4>(),
            d->hasv.as<unsigned char>(), d->V.as<float>(), d->cgV.as<float>(), ctx->stage[1].as<uint32_t>(),
            ctx->stage[2].as<int>(), ctx->stage[3].as<float>(), ctx->stage[4].as<float>(), d->counters.as<long long>());
  WH_CUDA(cudaMemcpyAsync(fea_cnt, ctx->stage[1].p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaMemcpyAsync(size, ctx->stage[2].p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaMemcpyAsync(w, ctx->stage[3].p, n * (dim + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaMemcpyAsync(cg, ctx->stage[4].p, n * (dim + 2) * 4, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  return WH_OK;
}

int wh_difacto_loss(whContext* ctx, const whLocalBlock* D, const float* model, size_t model_len, const int* model_siz,
                    size_t nkeys, int dim, float clip, float norm, double* prog, float* grad) {
  return wh_difacto_loss_dropout(ctx, D, model, model_len, model_siz, nkeys, dim, clip, norm, 0.f, nullptr, 0, prog, grad);
}

int wh_difacto_loss_dropout(whContext* ctx, const whLocalBlock* D, const float* model, size_t model_len, const int* model_siz,
                            size_t nkeys, int dim, float clip, float norm, float dropout, const int* rand_values,
                            size_t num_rand, double* prog, float* grad) {
  WH_REQUIRE(ctx && D && prog && (nkeys == 0 || (model_siz && (model || model_len == 0))), "wh_difacto_loss: NULL argument");
  WH_REQUIRE(dropout >= 0.f && dropout <= 1.f, "wh_difacto_loss: dropout %g outside [0, 1]", (double)dropout);
  if (dropout > 0.f && grad && dim > 0) {
    size_t need = 0;
    for (size_t i = 0; i < nkeys; ++i) if (model_siz[i] == dim + 1) need += (size_t)dim;
    WH_REQUIRE(need == 0 || (rand_values && num_rand >= need),
               "wh_difacto_loss: dropout needs the caller's next %zu rand() values (one per embedding-gradient float), %zu given", need, num_rand);
    if (need == 0) dropout = 0.f;
  }
  WH_REQUIRE(dim >= 0 && dim <= 256, "wh_difacto_loss: dim %d outside [0, 256]", dim);
  DeviceGuard g(ctx->device);
  // a throw-away DiFacto object supplies the scratch (no shard tables: dim-sized buffers only)
  whDifacto tmp;
  tmp.ctx = ctx; memset(&tmp.c, 0, sizeof(tmp.c)); tmp.c.dim = dim;
  int rc = WH_OK;
  auto body = [&]() -> int {
    WH_TRY(tmp.scal.reserve(4096)); WH_TRY(tmp.pinned.reserve(4096));
    WH_CUDA(cudaMemsetAsync(tmp.scal.p, 0, 4096, ctx->stream));
    tmp.ax.device = ctx->device; tmp.ax.num_sms = ctx->num_sms;
    WH_CUDA(cudaStreamCreateWithFlags(&tmp.ax.stream, cudaStreamNonBlocking));
    WH_TRY(tmp.ax.small.reserve(1 << 14));
    WH_CUDA(cudaEventCreateWithFlags(&tmp.ev_fork, cudaEventDisableTiming));
    WH_CUDA(cudaEventCreateWithFlags(&tmp.ev_auc, cudaEventDisableTiming));
    DevLocal dl;
    WH_TRY(stage_local(ctx, D, WH_HOST, &dl));
    for (int i = 0; i < 8; ++i) prog[i] = 0;
    if (dl.rows == 0) return WH_OK;
    WH_REQUIRE(dl.label != nullptr, "wh_difacto_loss: label is NULL");
    const size_t U = nkeys, sd = dim > 0 ? dim : 1;
    const float* d_model; const int* d_siz;
    WH_TRY(to_device(ctx, ctx->stage[4], model, model_len, WH_HOST, &d_model));
    WH_TRY(to_device(ctx, ctx->stage[5], model_siz, nkeys, WH_HOST, &d_siz));
    WH_TRY(ctx->stage[6].reserve((U ? U : 1) * 4));
    WH_TRY(tmp.w.reserve((U ? U : 1) * 4)); WH_TRY(tmp.has.reserve(U ? U : 1)); WH_TRY(tmp.Vs.reserve((U ? U : 1) * sd * 4));
    WH_TRY(ctx->scan_tmp.reserve(div_up(U ? U : 1, kScanTile) * sizeof(unsigned) + 256));
    unsigned* pos = ctx->stage[6].as<unsigned>();
    unsigned* total = tmp.scal.as<unsigned>() + 8;
    WH_TRY(scan_and_apply<unsigned>(ctx, SizVal{d_siz}, PosOut{pos}, U, ctx->scan_tmp.as<unsigned>(), total));
    WH_LAUNCH(dfc_unpack_kernel, (unsigned)div_up(U, 256), 256, 0, ctx->stream, U, dim, d_model, d_siz, pos,
              tmp.w.as<float>(), tmp.has.as<unsigned char>(), tmp.Vs.as<float>());
    DfcWork k;
    WH_TRY(dfc_work(&tmp, dl.rows, U, dl.nnz, &k));
    DfcDropout drop;
    if (dropout > 0.f && grad && dim > 0 && U) {
      WH_TRY(tmp.drop_stream.reserve(num_rand * 4)); WH_TRY(tmp.drop_pos.reserve(U * 4)); WH_TRY(tmp.counters.reserve(64));
      WH_CUDA(cudaMemcpyAsync(tmp.drop_stream.p, rand_values, num_rand * 4, cudaMemcpyHostToDevice, ctx->stream));
      WH_CUDA(cudaMemsetAsync(tmp.counters.p, 0, 64, ctx->stream));
      drop.dropout = dropout; drop.stream = tmp.drop_stream.as<int>(); drop.len = num_rand;
      drop.cursor = tmp.counters.as<long long>() + 4; drop.pos = tmp.drop_pos.as<unsigned>(); drop.total = tmp.scal.as<unsigned>() + 10;
    }
    WH_TRY(difacto_loss_launch(ctx, &tmp.ax, tmp.ev_fork, tmp.ev_auc, dl.rows, dl.nnz, dl.offset, dl.index, dl.value, dl.label,
                               U, (int)sd, tmp.w.as<float>(), tmp.has.as<unsigned char>(), tmp.Vs.as<float>(), dim > 0, clip,
                               norm, grad != nullptr, k, &drop));
    if (grad && model_len) {
      WH_TRY(ctx->stage[7].reserve(model_len * 4));
      WH_CUDA(cudaMemsetAsync(ctx->stage[7].p, 0, model_len * 4, ctx->stream));
      WH_LAUNCH(dfc_pack_kernel, (unsigned)div_up(U, 256), 256, 0, ctx->stream, U, dim, d_siz, pos, k.gw,
                tmp.has.as<unsigned char>(), k.G, ctx->stage[7].as<float>());
      WH_CUDA(cudaMemcpyAsync(grad, ctx->stage[7].p, model_len * 4, cudaMemcpyDeviceToHost, ctx->stream));
    }
    char* pin = tmp.pinned.as<char>();
    WH_CUDA(cudaMemcpyAsync(pin, k.evW, sizeof(EvalOut), cudaMemcpyDeviceToHost, ctx->stream));
    WH_CUDA(cudaMemcpyAsync(pin + 256, k.ev, sizeof(EvalOut), cudaMemcpyDeviceToHost, ctx->stream));
    WH_CUDA(cudaStreamSynchronize(ctx->stream));
    dfc_prog(reinterpret_cast<const EvalOut*>(pin), reinterpret_cast<const EvalOut*>(pin + 256), dl.rows, prog);
    return WH_OK;
  };
  rc = body();
  cudaStreamSynchronize(ctx->stream);
  if (tmp.ax.stream) { cudaStreamSynchronize(tmp.ax.stream); cudaStreamDestroy(tmp.ax.stream); }
  if (tmp.ev_fork) cudaEventDestroy(tmp.ev_fork);
  if (tmp.ev_auc) cudaEventDestroy(tmp.ev_auc);
  for (wh::DevBuf* b : {&tmp.w, &tmp.has, &tmp.Vs, &tmp.XV, &tmp.xxvv, &tmp.py, &tmp.dual, &tmp.gw, &tmp.G, &tmp.accG,
                        &tmp.accX, &tmp.scal, &tmp.vrows, &tmp.vidx, &tmp.vval, &tmp.klist, &tmp.ax.small, &tmp.ax.scan_tmp, &tmp.drop_stream, &tmp.drop_pos, &tmp.counters})
    b->release();
  for (auto& b : tmp.ax.sort_tmp) b.release();
  tmp.pinned.release();
  return rc;
}

}  // extern "C"
