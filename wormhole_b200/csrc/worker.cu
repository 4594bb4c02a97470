// The per-minibatch pipeline of one worker: AsgdWorker::ProcessMinibatch
// (learn/linear/async_sgd.h:240-288) with the server on the same GPU.
//
//   Localize (:246-247) -> ZPull (:287) -> Init = X w (:260) -> Evaluate (:261) ->
//   CalcGrad = X^T dual (:269) -> ZPush (:277-278)
//
// Minibatches in flight. The reference keeps up to max_concurrency minibatches in flight per
// worker (learn/solver/minibatch_solver.h:289-290,307-322). Here a worker owns `depth` lanes,
// each with its own CUDA stream and scratch; wh_worker_submit() enqueues a whole minibatch on
// the next lane without any host round trip (the number of unique keys stays on the device) and
// wh_worker_wait() collects its progress numbers. Lanes overlap freely except where the model
// is touched: the pull of minibatch t+1 waits (CUDA event) for the push of minibatch t, so
// every pull sees all earlier pushes -- the same numbers as strictly sequential processing
// (max_concurrency = 1, the configuration parity is judged in), while staging, localize and AUC
// of neighbouring minibatches run side by side. AUC, which nothing downstream depends on, runs on
// a side stream of the lane.
#include <stdlib.h>

#include <algorithm>
#include <chrono>

#include <vector>

#include "batch.cuh"
#include "direct.cuh"
#include "hot.cuh"
#include "loss.cuh"
#include "server.cuh"
#include "spmv.cuh"

namespace wh {
int spmv_trans_accumulate_launch(whContext* ctx, size_t rows, size_t nnz, const size_t* offset,
                                 const uint32_t* index, const float* value, const float* x,
                                 long long* acc, FixedScale* fs, unsigned copies, size_t stride, bool scale_ready = false);
int fixed_finish_launch(whContext* ctx, long long* acc, size_t n, const FixedScale* fs, float* y, float mul,
                        unsigned copies, size_t stride, const unsigned* d_n = nullptr, unsigned ystride = 1);
constexpr int kMaxDepth = 8;

struct Lane {
  whContext cx;              // main chain: own stream + directory / scan / accumulator scratch
  whContext ax;              // AUC side stream + its sort scratch
  cudaEvent_t ev_fork = nullptr, ev_auc = nullptr, ev_done = nullptr;
  DevBuf in_index, in_offset, in_value, in_label;   // staged host input
  DevBuf uniq, index, w, xw, dual, scal;
  char* pinned = nullptr;    // 512 B of the worker's pinned block
  DevBuf outbox;             // sharded jobs: {U, cuts | keys | grads}, mapped by every peer
  DevBuf hot;                // hot-key state of the lane (hot.cuh): present | rank | w, kHotSlots each
  DirectLane dx;             // accumulators of the shared-GPU path (direct.cuh)
  bool busy = false, train = false;
  size_t rows = 0;
  // state between wh_worker_begin and wh_worker_finish
  bool b_open = false;
  size_t b_nnz = 0, b_nuniq = 0;
  const size_t* b_offset = nullptr; const float* b_value = nullptr; const float* b_label = nullptr;
  unsigned long long ticket = 0;
};
}  // namespace wh

struct whWorker {
  whContext* ctx = nullptr;
  whServer* server = nullptr;
  int loss = WH_LOSS_LOGIT;
  uint64_t max_key = WH_NO_MAX_KEY;
  int depth = 1;
  wh::Lane lane[wh::kMaxDepth];
  wh::PinnedBuf pinned;
  // ev_pushed[t % ring]: completion of minibatch t's push (or, for evaluation, of its pull)
  cudaEvent_t ev_pushed[wh::kMaxDepth + 1] = {nullptr};
  // sharded jobs with owner-served pull: ev_served[t % ring] = every owner has served the pull of minibatch t (and
  // therefore finished applying minibatch t - 1): the lane outbox of minibatch t - 1 may be overwritten after it
  cudaEvent_t ev_served[wh::kMaxDepth + 1] = {nullptr};
  int max_delay = 0;                 // pull(t) waits for push(t - 1 - max_delay); pushes stay in order
  // sharded jobs: every rank's lane outboxes, [rank][lane] (own ones included); 0 capacity = not attached
  size_t outbox_cap = 0;
  unsigned char* peer_outbox[wh::kMaxShards][wh::kMaxDepth] = {{nullptr}};
  void* opened[wh::kMaxShards * wh::kMaxDepth] = {nullptr};
  unsigned long long next_ticket = 1, next_wait = 1;
  long long seen_new_w = 0;
  // hot keys (hot.cuh): the table image shared by the lanes; built from a census of the first eligible minibatch
  // ... and rebuilt every hot_refresh eligible minibatches (WH_HOT_REFRESH, default 2048, 0 = never), for streams whose
  // popular keys drift. Two image buffers alternate: kernels already enqueued keep the pointer they were launched with.
  wh::DevBuf hot_tags[2];
  int hot_cur = 0;                       // the buffer new minibatches use
  bool hot_enabled = true, hot_tried = false, hot_ready = false;
  int hot_keys = 0;
  long long hot_refresh = 2048, hot_since = 0;
  // worker and server on one GPU, hashed keys: keys address the shard directly (direct.cuh). WH_DIRECT=0 keeps the
  // localize -> pull -> forward/backward -> push chain of launches.
  bool dx_enabled = true, dx_tried = false;
  wh::DirectTable dx;
  long long dx_since = 0;
  double census_ms = 0;      // wall time of the last census (either table), for wh_worker_table_keys
  int census_count = 0;
};

using namespace wh;

static void lane_release(Lane& l) {
  for (DevBuf* b : {&l.in_index, &l.in_offset, &l.in_value, &l.in_label, &l.uniq, &l.index, &l.w, &l.xw, &l.dual,
                    &l.scal, &l.outbox, &l.hot, &l.dx.bitmap, &l.dx.part, &l.dx.part_touched, &l.dx.codes, &l.dx.touched, &l.cx.directory, &l.cx.scan_tmp, &l.cx.scan_state, &l.cx.acc, &l.cx.auc, &l.ax.auc, &l.cx.small, &l.ax.scan_tmp, &l.ax.small})
    b->release();
  for (auto& b : l.cx.sort_tmp) b.release();
  for (auto& b : l.ax.sort_tmp) b.release();
  for (auto& b : l.cx.stage) b.release();
  for (auto& b : l.ax.stage) b.release();
  for (auto e : l.cx.prof.ev) cudaEventDestroy(e);
  if (l.cx.side_ev) cudaEventDestroy(l.cx.side_ev);
  if (l.ev_fork) cudaEventDestroy(l.ev_fork);
  if (l.ev_auc) cudaEventDestroy(l.ev_auc);
  if (l.ev_done) cudaEventDestroy(l.ev_done);
  if (l.cx.stream && l.cx.own_stream) cudaStreamDestroy(l.cx.stream);
  if (l.ax.stream) cudaStreamDestroy(l.ax.stream);
}

static int lane_init(whWorker* w, int i) {
  Lane& l = w->lane[i];
  whContext* ctx = w->ctx;
  l.cx.device = l.ax.device = ctx->device;
  l.cx.num_sms = l.ax.num_sms = ctx->num_sms;
  l.cx.l2_bytes = l.ax.l2_bytes = ctx->l2_bytes;
  if (i == 0) { l.cx.stream = ctx->stream; l.cx.own_stream = false; }   // lane 0 rides the context's stream
  else {
    // the lanes' own streams carry the model-dependent chain: above the evaluation side streams (WH_PRIO=0: all alike)
    int least = 0, greatest = 0;
    WH_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    const char* e = getenv("WH_PRIO");
    const int prio = (e && atoi(e) == 0) ? least : greatest;
    WH_CUDA(cudaStreamCreateWithPriority(&l.cx.stream, cudaStreamNonBlocking, prio));
    l.cx.own_stream = true;
  }
  WH_CUDA(cudaStreamCreateWithFlags(&l.ax.stream, cudaStreamNonBlocking));
  l.ax.own_stream = true;
  WH_CUDA(cudaEventCreateWithFlags(&l.ev_fork, cudaEventDisableTiming));
  WH_CUDA(cudaEventCreateWithFlags(&l.ev_auc, cudaEventDisableTiming));
  WH_CUDA(cudaEventCreateWithFlags(&l.ev_done, cudaEventDisableTiming));
  WH_TRY(l.scal.reserve(4096));
  WH_TRY(l.hot.reserve((size_t)kHotSlots * (4 + 4 + 4 + 8)));
  WH_CUDA(cudaMemsetAsync(l.hot.p, 0, (size_t)kHotSlots * (4 + 4 + 4 + 8), l.cx.stream));
  WH_TRY(l.cx.small.reserve(1 << 14));
  WH_TRY(l.ax.small.reserve(1 << 14));
  l.pinned = w->pinned.as<char>() + 512 * i;
  // the AUC histogram of the lane's side context: allocated and zeroed here, so that the first fused kernel (which
  // adds to it from the main stream) finds it ready
  void* hist;
  WH_TRY(auc_scratch_get(&l.ax, &hist));
  WH_CUDA(cudaStreamSynchronize(l.ax.stream));
  return WH_OK;
}

static HotView lane_hot_view(const whWorker* w, Lane& l) {
  HotView v;
  v.tags = w->hot_tags[w->hot_cur].as<uint4>();
  v.present = l.hot.as<unsigned>();
  v.rank = v.present + kHotSlots;
  v.w = reinterpret_cast<float*>(v.rank + kHotSlots);
  return v;
}

// Census of one minibatch: localize it with counts, bring (key, count) to the host, pick the table. Runs once per
// worker, synchronously; its only product is the table image (a cache hint -- results never depend on it).
static int hot_census(whWorker* w, Lane& l, const LocInput& in, size_t ucap, bool for_direct) {
  whContext* cx = &l.cx;
  if (!l.cx.own_stream) l.cx.stream = w->ctx->stream;   // lane 0 rides the context's stream, whichever that is now (wh_set_stream)
  if (for_direct) w->dx_tried = true; else w->hot_tried = true;
  const auto t_begin = std::chrono::steady_clock::now();
  struct Timer { whWorker* w; std::chrono::steady_clock::time_point t0;
                 ~Timer() { w->census_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); w->census_count++; } } timer{w, t_begin};
  DevBuf keys, cnt, idx;
  struct Free { DevBuf *a, *b, *c; ~Free() { a->release(); b->release(); c->release(); } } guard{&keys, &cnt, &idx};
  WH_TRY(keys.reserve(ucap * 8)); WH_TRY(cnt.reserve(ucap * 4)); WH_TRY(idx.reserve(in.nnz * 4));
  unsigned* d_n = l.scal.as<unsigned>() + 16;
  WH_TRY(localize_direct(cx, in, keys.as<uint64_t>(), cnt.as<uint32_t>(), idx.as<uint32_t>(), d_n));
  unsigned U = 0;
  WH_CUDA(cudaMemcpyAsync(&U, d_n, 4, cudaMemcpyDeviceToHost, cx->stream));
  WH_CUDA(cudaStreamSynchronize(cx->stream));
  std::vector<uint64_t> hk(U);
  std::vector<uint32_t> hc(U), image(kHotSlots);
  if (U) {
    WH_CUDA(cudaMemcpy(hk.data(), keys.p, (size_t)U * 8, cudaMemcpyDeviceToHost));
    WH_CUDA(cudaMemcpy(hc.data(), cnt.p, (size_t)U * 4, cudaMemcpyDeviceToHost));
  }
  if (for_direct) {
    // the table is shared by the lanes and single-buffered: let the minibatches in flight finish first
    for (int i = 0; i < w->depth; ++i) WH_CUDA(cudaStreamSynchronize(w->lane[i].cx.stream));
    w->dx_since = 0;
    WH_TRY(direct_table_build(cx, &w->dx, hk.data(), hc.data(), U));
    w->dx.seen_version = ~0ull;   // the weight mirror has to be filled
    return WH_OK;
  }
  w->hot_since = 0;
  const int placed = hot_build_image(hk.data(), hc.data(), U, image.data());
  if (placed < 64) return WH_OK;                 // nothing repeats enough to be worth a (new) table
  const int next = w->hot_ready ? 1 - w->hot_cur : w->hot_cur;   // a refresh fills the buffer that is not in use
  WH_TRY(w->hot_tags[next].reserve((size_t)kHotSlots * 4));
  WH_CUDA(cudaMemcpy(w->hot_tags[next].p, image.data(), (size_t)kHotSlots * 4, cudaMemcpyHostToDevice));
  w->hot_cur = next;
  w->hot_keys = placed;
  w->hot_ready = true;
  return WH_OK;
}

extern "C" {

int wh_worker_create_ex(whContext* ctx, whServer* server, int loss, uint64_t max_key, int depth, whWorker** out) {
  WH_REQUIRE(ctx && out, "wh_worker_create: NULL argument");
  WH_REQUIRE(server == nullptr || server->ctx == ctx, "wh_worker_create: server lives on another context");
  WH_REQUIRE(loss == WH_LOSS_LOGIT || loss == WH_LOSS_SQUARE_HINGE || loss == WH_LOSS_SQUARE, "wh_worker_create: unknown loss type %d", loss);
  WH_REQUIRE(max_key != 0, "wh_worker_create: max_key must be positive");
  WH_REQUIRE(depth >= 1 && depth <= kMaxDepth, "wh_worker_create: depth %d outside [1, %d]", depth, kMaxDepth);
  DeviceGuard g(ctx->device);
  whWorker* w = new whWorker();
  w->ctx = ctx; w->server = server; w->loss = loss; w->max_key = max_key; w->depth = depth;
  { const char* e = getenv("WH_HOT"); w->hot_enabled = !(e && atoi(e) == 0); }   // WH_HOT=0: plain kernels only
  { const char* e = getenv("WH_HOT_REFRESH"); if (e) w->hot_refresh = atoll(e); }
  { const char* e = getenv("WH_DIRECT"); w->dx_enabled = !(e && atoi(e) == 0); }
  int r = w->pinned.reserve(512 * kMaxDepth);
  for (int i = 0; r == WH_OK && i < depth; ++i) r = lane_init(w, i);
  for (int i = 0; r == WH_OK && i <= kMaxDepth; ++i)
    if (cudaEventCreateWithFlags(&w->ev_pushed[i], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&w->ev_served[i], cudaEventDisableTiming) != cudaSuccess) r = WH_ERR_CUDA;
  if (r != WH_OK) { for (int i = 0; i < depth; ++i) lane_release(w->lane[i]); w->pinned.release(); delete w; return r; }
  *out = w;
  return WH_OK;
}

int wh_worker_create(whContext* ctx, whServer* server, int loss, uint64_t max_key, whWorker** out) {
  return wh_worker_create_ex(ctx, server, loss, max_key, 1, out);
}

int wh_worker_destroy(whWorker* w) {
  if (!w) return WH_OK;
  DeviceGuard g(w->ctx->device);
  for (int i = 0; i < w->depth; ++i) {
    cudaStreamSynchronize(w->lane[i].cx.stream);
    cudaStreamSynchronize(w->lane[i].ax.stream);
  }
  for (void* p : w->opened) if (p) cudaIpcCloseMemHandle(p);
  for (int i = 0; i < w->depth; ++i) lane_release(w->lane[i]);
  for (auto e : w->ev_pushed) if (e) cudaEventDestroy(e);
  for (auto e : w->ev_served) if (e) cudaEventDestroy(e);
  w->pinned.release();
  w->hot_tags[0].release(); w->hot_tags[1].release();
  w->dx.release();
  delete w;
  return WH_OK;
}

int wh_worker_wait(whWorker* w, uint64_t ticket, double* prog);

int wh_worker_table_keys(whWorker* w, uint32_t* keys, size_t cap, size_t* n, double* census_ms, int* censuses) {
  WH_REQUIRE(w && n, "wh_worker_table_keys: NULL argument");
  DeviceGuard g(w->ctx->device);
  *n = 0;
  if (census_ms) *census_ms = w->census_ms;
  if (censuses) *censuses = w->census_count;
  const void* src = nullptr; size_t slots = 0;
  if (w->dx.ready) { src = w->dx.tags.p; slots = kDxSlots; }
  else if (w->hot_ready) { src = w->hot_tags[w->hot_cur].p; slots = kHotSlots; }
  if (!src) return WH_OK;
  std::vector<uint32_t> img(slots);
  WH_CUDA(cudaMemcpy(img.data(), src, slots * 4, cudaMemcpyDeviceToHost));
  for (uint32_t k : img) if (k != 0xffffffffu) { if (keys && *n < cap) keys[*n] = k; ++*n; }
  return WH_OK;
}

int wh_worker_set_max_delay(whWorker* w, int max_delay) {
  WH_REQUIRE(w && max_delay >= 0 && max_delay < w->depth, "wh_worker_set_max_delay: needs 0 <= max_delay < depth");
  WH_REQUIRE(w->next_wait == w->next_ticket, "wh_worker_set_max_delay: minibatches are in flight");
  // sharded jobs: only with owner-side push (wh_worker_ipc_attach done), whose two barriers sit inside the push
  // phase and therefore stay in ticket order; every rank must choose the same value
  WH_REQUIRE(max_delay == 0 || !w->server || w->server->map.world == 1 || w->outbox_cap != 0,
             "wh_worker_set_max_delay: a sharded job needs wh_worker_ipc_attach first (push in turns keeps max_delay 0)");
  w->max_delay = max_delay;
  return WH_OK;
}

int wh_worker_submit(whWorker* w, const whRowBlock* blk, int mem, size_t nnz, int train, uint64_t* ticket) {
  WH_REQUIRE(w && blk && ticket, "wh_worker_submit: NULL argument");
  WH_REQUIRE(w->server != nullptr, "wh_worker_submit: this worker has no local server; use wh_worker_begin / wh_worker_finish");
  WH_REQUIRE(mem == WH_HOST || mem == WH_DEVICE, "wh_worker_submit: bad mem %d", mem);
  whContext* ctx = w->ctx;
  whServer* s = w->server;
  DeviceGuard g(ctx->device);
  const size_t rows = blk->size;
  WH_REQUIRE(rows > 0, "wh_worker_submit: empty minibatch");
  WH_REQUIRE(blk->offset && blk->index && blk->label, "wh_worker_submit: offset/index/label is NULL");
  Lane& l = w->lane[w->next_ticket % w->depth];
  if (l.busy) {  // the lane's previous minibatch has not been collected: results would be lost
    set_error("wh_worker_submit: %d minibatches already in flight; wh_worker_wait the oldest first", w->depth);
    return WH_ERR_INVALID;
  }
  whContext* cx = &l.cx;
  if (!l.cx.own_stream) l.cx.stream = w->ctx->stream;   // lane 0 rides the context's stream, whichever that is now (wh_set_stream)
  cx->prof.on = ctx->prof.on;
  size_t n = nnz;
  if (mem == WH_HOST) n = blk->offset[rows];
  else if (n == 0) {
    WH_CUDA(cudaMemcpyAsync(l.pinned + 384, blk->offset + rows, sizeof(size_t), cudaMemcpyDeviceToHost, cx->stream));
    WH_CUDA(cudaStreamSynchronize(cx->stream));
    n = *reinterpret_cast<size_t*>(l.pinned + 384);
  }
  WH_REQUIRE(n > 0 && n < (size_t)UINT32_MAX, "wh_worker_submit: %zu entries", n);
  const bool direct = direct_supported(w->max_key);

  // stage the minibatch
  const uint64_t* d_index; const size_t* d_offset; const float* d_value; const float* d_label;
  WH_TRY(to_device(cx, l.in_index, blk->index, n, mem, &d_index));
  WH_TRY(to_device(cx, l.in_offset, blk->offset, rows + 1, mem, &d_offset));
  WH_TRY(to_device(cx, l.in_value, blk->value, n, mem, &d_value));
  WH_TRY(to_device(cx, l.in_label, blk->label, rows, mem, &d_label));

  const size_t ucap = (size_t)std::min<uint64_t>(n, w->max_key);
  const bool outboxed = w->outbox_cap != 0 && s->map.world > 1;
  WH_REQUIRE(!outboxed || ucap <= w->outbox_cap, "wh_worker_submit: %zu possible unique keys exceed the outbox capacity %zu given to wh_worker_ipc_export",
             ucap, w->outbox_cap);
  if (!outboxed) WH_TRY(l.uniq.reserve(ucap * 8));
  WH_TRY(l.index.reserve(n * 4)); WH_TRY(l.w.reserve(ucap * 4));
  WH_TRY(l.xw.reserve(rows * 4)); WH_TRY(l.dual.reserve(rows * 4));
  // sharded jobs keep the key list, its count and the gradients in the lane's outbox, where the owners read them
  uint64_t* uniq = outboxed ? reinterpret_cast<uint64_t*>(l.outbox.as<unsigned char>() + 256) : l.uniq.as<uint64_t>();
  unsigned* d_nuniq = outboxed ? l.outbox.as<unsigned>() : l.scal.as<unsigned>();
  FixedScale* fs = reinterpret_cast<FixedScale*>(l.scal.as<char>() + 256);
  EvalOut* d_eval = reinterpret_cast<EvalOut*>(l.scal.as<char>() + 512);
  long long* acc = nullptr;
  if (train) WH_TRY(ensure_acc(cx, ucap, &acc));

  LocInput in{d_index, n, make_xform(w->max_key)};
  // Worker and server on one GPU, hashed key space, the shard covering all of it: keys address the shard directly,
  // two launches per minibatch (direct.cu)
  const bool dx_ok = w->dx_enabled && direct && s->map.world == 1 && s->hslots == 0 && s->key_begin == 0 &&
                     s->key_end >= w->max_key && n < (1ull << 31) && rows < (1ull << 31) &&
                     w->max_delay == 0 &&   // (the accumulators live in the shard: one minibatch adds at a time)
                     w->loss != WH_LOSS_SQUARE;   // (|dual| <= 1 sizes the fixed point up front; the square loss has no such bound)
  if (dx_ok) {
    const long long dx_refresh = w->hot_refresh > 0 ? std::max<long long>(w->hot_refresh, w->depth + 1) : 0;
    if (!w->dx_tried || (dx_refresh > 0 && ++w->dx_since >= dx_refresh)) WH_TRY(hot_census(w, l, in, ucap, true));
  }
  // The step pays a RED, a DRAM line and a bitmap bit for every entry whose key is NOT in the table; the chain of
  // launches below turns every key into a rank first and pays per unique key instead. KDD2012 shape (47 % of the
  // entries covered): 1.52 against 1.19 G examples/s in favour of the step; below 40 % (flat key distributions,
  // unmeasured) the chain is kept. WH_DX_MIN_COVERAGE overrides.
  double min_cov = 0.4;
  { const char* e = getenv("WH_DX_MIN_COVERAGE"); if (e) min_cov = atof(e); }
  if (dx_ok && w->dx.ready && (w->dx.coverage >= min_cov || getenv("WH_DX_TABLE"))) {
    const unsigned long long tk = w->next_ticket;
    const int ring = kMaxDepth + 1;
    WH_TRY(direct_lane_prepare(cx, &l.dx, w->max_key));
    // WH_DX_STAGED=1 (opt-in): with several minibatches in flight, the model-independent part of the step (ids -> codes)
    // is launched ahead of the wait for the previous minibatch's push, to overlap with that minibatch's step. Measured at
    // the Criteo shape: the step kernel shrinks from 65 to 54 us, but the stage kernel's 30 us only partly hide beside
    // it (both are bound by the same memory pipes): 91.7 us per step against 89.2 for the single launch.
    int staged_env = 0;
    { const char* e = getenv("WH_DX_STAGED"); staged_env = e ? atoi(e) : 0; }
    const bool staged = staged_env != 0 && w->depth > 1;
    DirectStep st;
    st.loss = w->loss; st.train = train != 0; st.rows = rows; st.nnz = n;
    st.offset = d_offset; st.index = d_index; st.value = d_value; st.label = d_label; st.xf = in.xf;
    st.shard = s->view().e; st.codes = nullptr;
    if (staged) { WH_TRY(prof_begin(cx)); WH_TRY(direct_stage_launch(cx, st, w->dx, l.dx)); st.codes = l.dx.codes.as<uint32_t>(); WH_PROF(cx, "stage_codes"); }
    if (tk > 1ull + (unsigned)w->max_delay) WH_CUDA(cudaStreamWaitEvent(cx->stream, w->ev_pushed[(tk - 1 - w->max_delay) % ring], 0));
    if (w->dx.seen_version != s->version) {
      // the shard was changed by someone else (wh_server_push / set_state / load, another worker): refill the mirror
      WH_CUDA(cudaStreamSynchronize(ctx->stream));
      for (int i = 0; i < w->depth; ++i) WH_CUDA(cudaStreamSynchronize(w->lane[i].cx.stream));
      WH_TRY(direct_table_refresh_w(cx, &w->dx, s->view().e));
      w->dx.seen_version = s->version;
    }
    if (!staged) WH_TRY(prof_begin(cx));
    if (d_value && train) { WH_TRY(fixed_prepare_unit_x(cx, fs, d_value, n)); WH_PROF(cx, "value_scale"); }
    void* hist;
    WH_TRY(auc_scratch_get(&l.ax, &hist));
    unsigned nparts = 0;
    WH_TRY(cx->stage[7].reserve((size_t)kDxMaxBlocks * 16 + 64));   // (a block per SM and wave, direct.cu)
    double* po = cx->stage[7].as<double>();
    unsigned* pc = reinterpret_cast<unsigned*>(po + kDxMaxBlocks);
    st.xw = l.xw.as<float>(); st.part_objv = po; st.part_correct = pc;
    st.auc_hist = auc_hist_ptr(hist); st.fs = fs; st.counters = s->counters.as<long long>();
    WH_TRY(direct_fused_launch(cx, st, w->dx, l.dx, &nparts));
    WH_PROF(cx, "fused_direct");
    // evaluation partials + AUC on the side stream (nothing downstream of them but the progress numbers)
    WH_CUDA(cudaEventRecord(l.ev_fork, cx->stream));
    WH_CUDA(cudaStreamWaitEvent(l.ax.stream, l.ev_fork, 0));
    WH_TRY(auc_count_launch(&l.ax, rows, d_label, l.xw.as<float>(), d_eval, true, nparts, po, pc));
    WH_CUDA(cudaEventRecord(l.ev_auc, l.ax.stream));
    if (train) {
      if (w->max_delay > 0 && tk > 1) WH_CUDA(cudaStreamWaitEvent(cx->stream, w->ev_pushed[(tk - 1) % ring], 0));
      WH_TRY(direct_push_launch(s, cx->stream, w->dx, l.dx, fs, w->loss == WH_LOSS_SQUARE_HINGE ? -2.f : 1.f, ctx->num_sms));
      w->dx.seen_version = s->version;
      WH_PROF(cx, "push");
    }
    WH_CUDA(cudaMemcpyAsync(l.pinned + 256, s->counters.p, 24, cudaMemcpyDeviceToHost, cx->stream));
    WH_CUDA(cudaEventRecord(w->ev_pushed[tk % ring], cx->stream));
    WH_CUDA(cudaStreamWaitEvent(cx->stream, l.ev_auc, 0));
    if (cx->prof.on) WH_TRY(prof_mark(cx, "auc_tail"));
    WH_CUDA(cudaMemcpyAsync(l.pinned, d_eval, sizeof(EvalOut), cudaMemcpyDeviceToHost, cx->stream));
    WH_CUDA(cudaEventRecord(l.ev_done, cx->stream));
    l.busy = true; l.train = train != 0; l.rows = rows;
    l.ticket = w->next_ticket++;
    *ticket = l.ticket;
    return WH_OK;
  }
  // hot keys (hot.cuh): binary features in a direct-addressed key space
  const bool hot_ok = w->hot_enabled && direct && d_value == nullptr && w->loss != WH_LOSS_SQUARE;
  // (a refresh overwrites the image that was current two refreshes ago: the interval must exceed the minibatches in flight)
  const long long refresh = w->hot_refresh > 0 ? std::max<long long>(w->hot_refresh, w->depth + 1) : 0;
  if (hot_ok && (!w->hot_tried || (refresh > 0 && ++w->hot_since >= refresh))) WH_TRY(hot_census(w, l, in, ucap, false));
  const bool hot = hot_ok && w->hot_ready;
  const bool fused = hot && train;    // forward + backward in one launch, which is also the rank pass of the localizer
  const HotView hv = lane_hot_view(w, l);
  // Accumulator replicas exist to spread the atomics on popular keys; the fused kernel adds those up in shared memory,
  // so one replica is enough there (measured: push 20.6 -> 16 us, step 121 -> 112 us with three minibatches in flight)
  const unsigned copies = fused ? 1u : acc_copies();

  const unsigned long long tk = w->next_ticket;       // this minibatch's ticket (1, 2, ...)
  const int ring = kMaxDepth + 1;
  static int serve = -1;   // WH_SERVE=0: every worker gathers its weights itself over NVLink (the older pull)
  if (serve < 0) { const char* e = getenv("WH_SERVE"); serve = e ? atoi(e) : 1; }
  const bool served = outboxed && serve != 0;
  // this lane's outbox still holds the keys and gradients of minibatch tk - depth, which the owners read while they
  // apply it; "every owner has served pull tk - depth + 1" says they are done with it
  if (served && w->depth > 1 && tk > (unsigned long long)w->depth)
    WH_CUDA(cudaStreamWaitEvent(cx->stream, w->ev_served[(tk - w->depth + 1) % ring], 0));
  cx->avoid_coop = outboxed && w->depth > 1;
  // With the table of frequent keys the step's kernel leaves its fixed-point sums in the outbox itself and the owners
  // convert while they apply (one launch less between push and pull): zero the words of the lane's previous minibatch,
  // and tell the owners how to read this one's
  const bool ob_fixed = outboxed && fused;
  if (outboxed) {
    WH_TRY(outbox_clear_launch(s, cx->stream, l.outbox.as<unsigned char>(), w->outbox_cap));
    OutboxScale* hs = reinterpret_cast<OutboxScale*>(l.pinned + 448);
    float sc_unused;
    hs->inv = 0; hs->mul = w->loss == WH_LOSS_SQUARE_HINGE ? -2.f : 1.f; hs->fixed = ob_fixed ? 1u : 0u;
    if (ob_fixed) unit_terms_scale(n, &sc_unused, &hs->inv);
    WH_CUDA(cudaMemcpyAsync(l.outbox.as<unsigned char>() + kOutboxScaleAt, hs, sizeof(OutboxScale), cudaMemcpyHostToDevice, cx->stream));
    if (ob_fixed) acc = reinterpret_cast<long long*>(l.outbox.as<unsigned char>() + outbox_grads_at(w->outbox_cap));
  }

  // local dense shard: Handle::Pull rides in the scan that emits the unique keys (after the push it has to see)
  const bool emit_pull = fused && s->map.world == 1 && s->hslots == 0 && pull_at_emit_supported(cx, w->max_key);
  PullAtEmit pe;
  if (emit_pull) {
    pe.table = s->view().e; pe.algo = s->algo; pe.key_begin = s->key_begin; pe.key_end = s->key_end;
    pe.w = l.w.as<float>(); pe.counters = s->counters.as<long long>();
    if (tk > 1ull + (unsigned)w->max_delay) pe.after = w->ev_pushed[(tk - 1 - w->max_delay) % ring];
  }

  WH_TRY(prof_begin(cx));
  // Localize
  if (hot) WH_TRY(localize_direct_hot(cx, in, hv, uniq, l.index.as<uint32_t>(), d_nuniq, !fused, emit_pull ? &pe : nullptr));
  else if (direct) WH_TRY(localize_direct(cx, in, uniq, nullptr, l.index.as<uint32_t>(), d_nuniq));
  else WH_TRY(localize_sort(cx, in, uniq, nullptr, l.index.as<uint32_t>(), d_nuniq));
  if (outboxed) WH_TRY(outbox_cuts_launch(s, cx->stream, l.outbox.as<unsigned char>()));
  // ZPull: one Handle::Pull per unique key, after every earlier push
  if (served) WH_TRY(server_barrier_launch(s, cx->stream, 1));   // every rank's key list is posted (off the critical cycle)
  if (g_timeline) WH_PROF(cx, "cuts+b1");
  // bounded staleness (the reference's max_concurrency > 1 lets a worker pull before its earlier pushes have
  // landed): with max_delay = d the pull of minibatch t waits for the push of minibatch t - 1 - d only
  if (!emit_pull && tk > 1ull + (unsigned)w->max_delay) WH_CUDA(cudaStreamWaitEvent(cx->stream, w->ev_pushed[(tk - 1 - w->max_delay) % ring], 0));
  const bool sharded = s->map.world > 1;
  float* pulled = l.w.as<float>();
  if (served) {
    // owner-served pull: each owner, after its own earlier pushes (the event above), answers the slices of all
    // ranks' key lists that it owns and writes the weights into their inboxes; "all served" completes the pull
    const int lane_idx = (int)(&l - w->lane);
    unsigned char* boxes[kMaxShards];
    for (int src = 0; src < s->map.world; ++src) boxes[src] = w->peer_outbox[src][lane_idx];
    if (g_timeline) WH_PROF(cx, "wait_push");
    WH_TRY(server_serve_all_launch(s, cx->stream, boxes, w->outbox_cap));
    if (g_timeline) WH_PROF(cx, "serve");
    WH_TRY(server_barrier_launch(s, cx->stream, 2));
    WH_CUDA(cudaEventRecord(w->ev_served[tk % ring], cx->stream));
    pulled = reinterpret_cast<float*>(l.outbox.as<unsigned char>() + outbox_inbox_at(w->outbox_cap));
  } else if (sharded) {
    // gather straight from the GPUs that own the keys; nobody may start pushing before everyone has pulled
    WH_TRY(server_pull_sharded_launch(s, cx->stream, uniq, ucap, d_nuniq, l.w.as<float>()));
    if (!outboxed || !train) WH_TRY(server_barrier_launch(s, cx->stream));
  } else if (!emit_pull) {
    WH_TRY(server_pull_launch(s, cx->stream, uniq, ucap, d_nuniq, l.w.as<float>()));
  }
  WH_PROF(cx, "pull");
  // Init (Xw), the per-example part of Evaluate, and the dual of CalcGrad: one pass over the rows. With
  // binary features |dual| <= 1 bounds every term of X^T dual, so the fixed-point scale is known up front.
  const bool scale_ready = d_value == nullptr && w->loss != WH_LOSS_SQUARE;
  if (hot && !fused) WH_TRY(hot_w_launch(cx, hv, pulled));   // evaluation: the slot-indexed weights for spmv_loss
  if (fused) {
    // ... and with the hot-key table the transposed product rides in the same launch (spmv.cu)
    WH_TRY(spmv_fused_hot_launch(cx, w->loss, rows, n, d_offset, l.index.as<uint32_t>(), d_label, pulled, hv,
                                 l.xw.as<float>(), nullptr, d_eval, fs, acc, copies, ucap,
                                 cx->directory.p, &l.ax));
    WH_PROF(cx, "spmv_fused");
  } else {
    WH_TRY(spmv_loss_launch(cx, w->loss, rows, n, d_offset, l.index.as<uint32_t>(), d_value, d_label, pulled,
                            l.xw.as<float>(), train ? l.dual.as<float>() : nullptr, d_eval,
                            (train && scale_ready) ? fs : nullptr, hot ? hv.w : nullptr));
    WH_PROF(cx, "spmv_loss");
  }
  // AUC on the side stream (nothing downstream of it but the progress numbers)
  WH_CUDA(cudaEventRecord(l.ev_fork, cx->stream));
  WH_CUDA(cudaStreamWaitEvent(l.ax.stream, l.ev_fork, 0));
  WH_TRY(auc_launch(&l.ax, rows, d_label, l.xw.as<float>(), d_eval, sharded && w->depth > 1));
  // the fused kernel was the last reader of the localizer's directory: zero it for the lane's next minibatch, off the chain
  if (fused) WH_CUDA(cudaMemsetAsync(cx->directory.p, 0, div_up(w->max_key, 32) * sizeof(uint2), l.ax.stream));
  WH_CUDA(cudaEventRecord(l.ev_auc, l.ax.stream));
  if (train) {
    // CalcGrad: X^T dual into the fixed-point accumulator; ZPush: convert + update in one pass
    if (!fused) {
      WH_TRY(spmv_trans_accumulate_launch(cx, rows, n, d_offset, l.index.as<uint32_t>(), d_value, l.dual.as<float>(), acc, fs, copies, ucap, scale_ready));
      WH_PROF(cx, "spmv_trans");
    }
    const float mul = w->loss == WH_LOSS_SQUARE_HINGE ? -2.f : 1.f;
    // pushes are applied one message at a time, in ticket order (two push kernels on one key would tear an entry)
    if (w->max_delay > 0 && tk > 1) WH_CUDA(cudaStreamWaitEvent(cx->stream, w->ev_pushed[(tk - 1) % ring], 0));
    if (outboxed) {
      // owners apply: gradients go to the outbox as fp32; once every rank has pulled and filled its outbox
      // (barrier), each owner walks the source ranks in ascending order -- one push message each -- reading its
      // slice straight from the source GPU; a second barrier keeps outboxes and tables stable until all are done
      if (!ob_fixed) {
        float* ob_grads = reinterpret_cast<float*>(l.outbox.as<unsigned char>() + outbox_grads_at(w->outbox_cap));
        WH_TRY(fixed_finish_launch(cx, acc, ucap, fs, ob_grads, mul, copies, ucap, d_nuniq, 2));   // fp32 in the low half of each word
        if (g_timeline) WH_PROF(cx, "fixed_finish");
      }
      const int lane_idx = (int)(&l - w->lane);
      static int apply_all = -1;   // WH_APPLY_ALL=0: one launch per source (the older path)
      if (apply_all < 0) { const char* e = getenv("WH_APPLY_ALL"); apply_all = e ? atoi(e) : 1; }
      // every rank's gradients are posted (with the owner-served pull: a barrier class of its own, see server.cuh)
      WH_TRY(server_barrier_launch(s, cx->stream, served ? 3 : 0));
      if (g_timeline) WH_PROF(cx, "b3");
      if (apply_all) {
        // the `world` messages in ascending source order in one cooperative launch
        unsigned char* boxes[kMaxShards];
        for (int src = 0; src < s->map.world; ++src) boxes[src] = w->peer_outbox[src][lane_idx];
        WH_TRY(server_apply_all_launch(s, cx->stream, boxes, w->outbox_cap));
      } else {
        for (int src = 0; src < s->map.world; ++src)
          WH_TRY(server_apply_from_launch(s, cx->stream, w->peer_outbox[src][lane_idx], w->outbox_cap));
      }
      // Nobody may pull from, or refill an outbox of, a step that some owner is still applying. With the
      // owner-served pull both follow from stream order -- an owner serves pull t+1 after its own apply t, and the
      // outbox is refilled only after "all served t+1" -- so no trailing barrier; one lane has no such slack.
      if (!served || w->depth == 1) WH_TRY(server_barrier_launch(s, cx->stream));
      s->steps++;
    } else if (sharded) {
      for (int turn = 0; turn < s->map.world; ++turn) {   // turn k: this rank is the only writer of shard (rank+k)%world
        WH_TRY(server_push_turn_launch(s, cx->stream, turn, uniq, ucap, d_nuniq, acc, fs, mul, copies, ucap));
        WH_TRY(server_barrier_launch(s, cx->stream));
      }
      s->steps++;
    } else {
      WH_TRY(server_push_acc_launch(s, cx->stream, uniq, ucap, d_nuniq, acc, fs, mul, copies, ucap));
    }
    WH_PROF(cx, "push");
  }
  // (the next minibatch's pull waits for this event: nothing but the push itself may sit in front of it)
  WH_CUDA(cudaEventRecord(w->ev_pushed[tk % ring], cx->stream));
  WH_CUDA(cudaMemcpyAsync(l.pinned + 256, s->counters.p, 24, cudaMemcpyDeviceToHost, cx->stream));
  // join the AUC branch, then the progress read-back
  WH_CUDA(cudaStreamWaitEvent(cx->stream, l.ev_auc, 0));
  if (cx->prof.on) WH_TRY(prof_mark(cx, "auc_tail"));
  WH_CUDA(cudaMemcpyAsync(l.pinned, d_eval, sizeof(EvalOut), cudaMemcpyDeviceToHost, cx->stream));
  WH_CUDA(cudaEventRecord(l.ev_done, cx->stream));
  l.busy = true; l.train = train != 0; l.rows = rows;
  l.ticket = w->next_ticket++;
  *ticket = l.ticket;
  return WH_OK;
}

// A step of a sharded job in which THIS rank has no minibatch while others still do (the ranks' file parts hold
// different numbers of minibatches, the barriers between the GPUs are collective): post an empty key list, serve the
// others' pulls, apply their pushes. No push message of this rank reaches any shard.
int wh_worker_submit_idle(whWorker* w, int train, uint64_t* ticket) {
  WH_REQUIRE(w && ticket && w->server, "wh_worker_submit_idle: NULL argument");
  whContext* ctx = w->ctx;
  whServer* s = w->server;
  WH_REQUIRE(s->map.world > 1 && w->outbox_cap != 0, "wh_worker_submit_idle: only for sharded jobs with outboxes (wh_worker_ipc_attach)");
  DeviceGuard g(ctx->device);
  Lane& l = w->lane[w->next_ticket % w->depth];
  if (l.busy) {
    set_error("wh_worker_submit_idle: %d minibatches already in flight; wh_worker_wait the oldest first", w->depth);
    return WH_ERR_INVALID;
  }
  whContext* cx = &l.cx;
  if (!l.cx.own_stream) l.cx.stream = w->ctx->stream;
  const unsigned long long tk = w->next_ticket;
  const int ring = kMaxDepth + 1;
  const char* e = getenv("WH_SERVE");
  const bool served = !e || atoi(e) != 0;
  const int lane_idx = (int)(&l - w->lane);
  EvalOut* d_eval = reinterpret_cast<EvalOut*>(l.scal.as<char>() + 512);
  unsigned char* boxes[kMaxShards];
  for (int src = 0; src < s->map.world; ++src) boxes[src] = w->peer_outbox[src][lane_idx];
  if (served && w->depth > 1 && tk > (unsigned long long)w->depth)
    WH_CUDA(cudaStreamWaitEvent(cx->stream, w->ev_served[(tk - w->depth + 1) % ring], 0));
  WH_TRY(outbox_clear_launch(s, cx->stream, l.outbox.as<unsigned char>(), w->outbox_cap));
  WH_CUDA(cudaMemsetAsync(l.outbox.p, 0, 256, cx->stream));      // U = 0, every cut = 0
  WH_CUDA(cudaMemsetAsync(d_eval, 0, sizeof(EvalOut), cx->stream));
  if (served) WH_TRY(server_barrier_launch(s, cx->stream, 1));
  if (tk > 1ull + (unsigned)w->max_delay) WH_CUDA(cudaStreamWaitEvent(cx->stream, w->ev_pushed[(tk - 1 - w->max_delay) % ring], 0));
  if (served) {
    WH_TRY(server_serve_all_launch(s, cx->stream, boxes, w->outbox_cap));
    WH_TRY(server_barrier_launch(s, cx->stream, 2));
    WH_CUDA(cudaEventRecord(w->ev_served[tk % ring], cx->stream));
  } else if (!train) {
    WH_TRY(server_barrier_launch(s, cx->stream));
  }
  if (train) {
    if (w->max_delay > 0 && tk > 1) WH_CUDA(cudaStreamWaitEvent(cx->stream, w->ev_pushed[(tk - 1) % ring], 0));
    WH_TRY(server_barrier_launch(s, cx->stream, served ? 3 : 0));
    WH_TRY(server_apply_all_launch(s, cx->stream, boxes, w->outbox_cap));
    if (!served || w->depth == 1) WH_TRY(server_barrier_launch(s, cx->stream));
    s->steps++;
  }
  WH_CUDA(cudaEventRecord(w->ev_pushed[tk % ring], cx->stream));
  WH_CUDA(cudaMemcpyAsync(l.pinned + 256, s->counters.p, 24, cudaMemcpyDeviceToHost, cx->stream));
  WH_CUDA(cudaMemcpyAsync(l.pinned, d_eval, sizeof(EvalOut), cudaMemcpyDeviceToHost, cx->stream));
  WH_CUDA(cudaEventRecord(l.ev_done, cx->stream));
  l.busy = true; l.train = train != 0; l.rows = 0;
  l.ticket = w->next_ticket++;
  *ticket = l.ticket;
  return WH_OK;
}

int wh_server_set_active_sources(whServer* s, uint64_t mask) {
  WH_REQUIRE(s, "wh_server_set_active_sources: NULL argument");
  s->active_sources = mask;
  return WH_OK;
}

int wh_worker_wait(whWorker* w, uint64_t ticket, double* prog) {
  WH_REQUIRE(w && prog, "wh_worker_wait: NULL argument");
  WH_REQUIRE(ticket == w->next_wait, "wh_worker_wait: tickets are collected in submission order (expected %llu, got %llu)",
             (unsigned long long)w->next_wait, (unsigned long long)ticket);
  whContext* ctx = w->ctx;
  whServer* s = w->server;
  DeviceGuard g(ctx->device);
  Lane& l = w->lane[ticket % w->depth];
  WH_REQUIRE(l.busy && l.ticket == ticket, "wh_worker_wait: ticket %llu is not in flight", (unsigned long long)ticket);
  WH_CUDA(cudaEventSynchronize(l.ev_done));
  l.busy = false;
  w->next_wait++;
  if (l.cx.prof.on) {
    WH_TRY(prof_collect(&l.cx));
    for (auto& a : l.cx.prof.acc) {   // fold the lane's numbers into the context's
      Profiler::Acc* t = nullptr;
      for (auto& x : ctx->prof.acc) if (x.name == a.name) { t = &x; break; }
      if (!t) { ctx->prof.acc.push_back(Profiler::Acc()); t = &ctx->prof.acc.back(); t->name = a.name; }
      t->ms += a.ms; t->count += a.count;
    }
    l.cx.prof.acc.clear();
  }
  const long long* cnt = reinterpret_cast<const long long*>(l.pinned + 256);
  if (cnt[2]) {
    WH_CUDA(cudaMemsetAsync(s->counters.as<long long>() + 2, 0, 8, l.cx.stream));
    set_error("wh_worker: a barrier between the GPUs of this job timed out (%lld times): a rank is missing or behind", cnt[2]);
    return WH_ERR_CUDA;
  }
  if (cnt[1]) {
    WH_CUDA(cudaMemsetAsync(s->counters.as<long long>() + 1, 0, 8, l.cx.stream));
    if (s->hslots)
      set_error("wh_worker: %lld keys found no slot: the server's open-addressing table of %llu slots is full (give a larger "
                "hash_slots), or they fall outside its range [%llu, %llu)", cnt[1], (unsigned long long)s->hslots,
                (unsigned long long)s->key_begin, (unsigned long long)s->key_end);
    else
      set_error("wh_worker: %lld keys fall outside the server's range [%llu, %llu)", cnt[1],
                (unsigned long long)s->key_begin, (unsigned long long)s->key_end);
    return WH_ERR_INVALID;
  }
  EvalHost h = eval_finish(*reinterpret_cast<const EvalOut*>(l.pinned), l.rows);
  prog[0] = h.objv; prog[1] = h.acc; prog[2] = h.auc; prog[3] = 1; prog[4] = (double)l.rows;
  if (l.rows == 0) prog[0] = prog[1] = prog[2] = prog[3] = 0;   // an idle step (wh_worker_submit_idle): nothing was evaluated
  prog[5] = (double)(cnt[0] - w->seen_new_w);
  w->seen_new_w = cnt[0];
  return WH_OK;
}

__global__ void predict_kernel(size_t n, float* xw, int prob) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  // learn/linear/loss.h:59: 1.0 / (1.0 + exp(-p)) with p a float: fp32 exp, the rest in double, printed as text
  if (i < n && prob) xw[i] = (float)(1.0 / (1.0 + (double)exp_rn(-xw[i])));
}

int wh_worker_predict(whWorker* w, const whRowBlock* blk, int mem, int prob, float* out) {
  WH_REQUIRE(w && blk && out, "wh_worker_predict: NULL argument");
  WH_REQUIRE(w->server && w->server->map.world == 1, "wh_worker_predict: needs a local, unsharded server");
  WH_REQUIRE(w->next_wait == w->next_ticket, "wh_worker_predict: minibatches are still in flight");
  whContext* ctx = w->ctx;
  DeviceGuard g(ctx->device);
  const size_t rows = blk->size;
  if (rows == 0) return WH_OK;
  WH_REQUIRE(blk->offset && blk->index, "wh_worker_predict: offset/index is NULL");
  Lane& l = w->lane[0];
  whContext* cx = &l.cx;
  if (!l.cx.own_stream) l.cx.stream = w->ctx->stream;   // lane 0 rides the context's stream, whichever that is now (wh_set_stream)
  size_t n;
  if (mem == WH_HOST) n = blk->offset[rows];
  else {
    WH_CUDA(cudaMemcpyAsync(l.pinned + 384, blk->offset + rows, sizeof(size_t), cudaMemcpyDeviceToHost, cx->stream));
    WH_CUDA(cudaStreamSynchronize(cx->stream));
    n = *reinterpret_cast<size_t*>(l.pinned + 384);
  }
  WH_TRY(l.xw.reserve(rows * 4));
  WH_CUDA(cudaMemsetAsync(l.xw.p, 0, rows * 4, cx->stream));
  if (n > 0) {
    const uint64_t* d_index; const size_t* d_offset; const float* d_value;
    WH_TRY(to_device(cx, l.in_index, blk->index, n, mem, &d_index));
    WH_TRY(to_device(cx, l.in_offset, blk->offset, rows + 1, mem, &d_offset));
    WH_TRY(to_device(cx, l.in_value, blk->value, n, mem, &d_value));
    const size_t ucap = (size_t)std::min<uint64_t>(n, w->max_key);
    WH_TRY(l.uniq.reserve(ucap * 8)); WH_TRY(l.index.reserve(n * 4)); WH_TRY(l.w.reserve(ucap * 4));
    unsigned* d_nuniq = l.scal.as<unsigned>();
    LocInput in{d_index, n, make_xform(w->max_key)};
    if (direct_supported(w->max_key)) WH_TRY(localize_direct(cx, in, l.uniq.as<uint64_t>(), nullptr, l.index.as<uint32_t>(), d_nuniq));
    else WH_TRY(localize_sort(cx, in, l.uniq.as<uint64_t>(), nullptr, l.index.as<uint32_t>(), d_nuniq));
    WH_TRY(server_pull_launch(w->server, cx->stream, l.uniq.as<uint64_t>(), ucap, d_nuniq, l.w.as<float>()));
    WH_TRY(spmv_times_launch(cx, rows, d_offset, l.index.as<uint32_t>(), d_value, l.w.as<float>(), l.xw.as<float>()));
  }
  WH_LAUNCH(predict_kernel, (unsigned)div_up(rows, 256), 256, 0, cx->stream, rows, l.xw.as<float>(), prob);
  WH_CUDA(cudaMemcpyAsync(out, l.xw.p, rows * 4, cudaMemcpyDeviceToHost, cx->stream));
  WH_CUDA(cudaStreamSynchronize(cx->stream));
  return WH_OK;
}

// ---- outboxes of a sharded job -----------------------------------------------------------------------

int wh_worker_ipc_export(whWorker* w, size_t max_uniq, void* handles) {
  WH_REQUIRE(w && handles && max_uniq > 0, "wh_worker_ipc_export: bad argument");
  WH_REQUIRE(w->server && w->outbox_cap == 0, "wh_worker_ipc_export: needs a worker with a local server, once");
  DeviceGuard g(w->ctx->device);
  cudaIpcMemHandle_t* h = reinterpret_cast<cudaIpcMemHandle_t*>(handles);
  for (int i = 0; i < w->depth; ++i) {
    Lane& l = w->lane[i];
    WH_TRY(l.outbox.reserve(outbox_bytes(max_uniq)));   // header | keys | gradients | inbox of pulled weights (server.cuh)
    WH_CUDA(cudaMemsetAsync(l.outbox.p, 0, 256, w->ctx->stream));
    WH_CUDA(cudaIpcGetMemHandle(&h[i], l.outbox.p));
  }
  WH_CUDA(cudaStreamSynchronize(w->ctx->stream));
  w->outbox_cap = max_uniq;
  return WH_OK;
}

int wh_worker_ipc_attach(whWorker* w, const void* handles) {
  WH_REQUIRE(w && handles && w->server, "wh_worker_ipc_attach: bad argument");
  WH_REQUIRE(w->outbox_cap != 0, "wh_worker_ipc_attach: call wh_worker_ipc_export first");
  const ShardMap& m = w->server->map;
  WH_REQUIRE(m.world > 1, "wh_worker_ipc_attach: attach the server to its peers first (wh_server_ipc_attach)");
  DeviceGuard g(w->ctx->device);
  const cudaIpcMemHandle_t* h = reinterpret_cast<const cudaIpcMemHandle_t*>(handles);
  for (int r = 0; r < m.world; ++r)
    for (int i = 0; i < w->depth; ++i) {
      if (r == m.rank) { w->peer_outbox[r][i] = w->lane[i].outbox.as<unsigned char>(); continue; }
      void* p = nullptr;
      WH_CUDA(cudaIpcOpenMemHandle(&p, h[r * w->depth + i], cudaIpcMemLazyEnablePeerAccess));
      w->opened[r * kMaxDepth + i] = p;
      w->peer_outbox[r][i] = static_cast<unsigned char*>(p);
    }
  return WH_OK;
}

// ---- the two halves around ZPull, for servers that live elsewhere (key-range shards on other GPUs)

int wh_worker_begin(whWorker* w, const whRowBlock* blk, int mem, size_t nnz, const uint64_t** d_uniq, size_t* num_uniq) {
  WH_REQUIRE(w && blk && d_uniq && num_uniq, "wh_worker_begin: NULL argument");
  WH_REQUIRE(mem == WH_HOST || mem == WH_DEVICE, "wh_worker_begin: bad mem %d", mem);
  WH_REQUIRE(w->next_wait == w->next_ticket, "wh_worker_begin: minibatches submitted earlier are still in flight");
  whContext* ctx = w->ctx;
  DeviceGuard g(ctx->device);
  const size_t rows = blk->size;
  WH_REQUIRE(rows > 0 && blk->offset && blk->index && blk->label, "wh_worker_begin: empty block or NULL offset/index/label");
  Lane& l = w->lane[0];
  whContext* cx = &l.cx;
  if (!l.cx.own_stream) l.cx.stream = w->ctx->stream;   // lane 0 rides the context's stream, whichever that is now (wh_set_stream)
  cx->prof.on = ctx->prof.on;
  size_t n = nnz;
  if (mem == WH_HOST) n = blk->offset[rows];
  else if (n == 0) {
    WH_CUDA(cudaMemcpyAsync(l.pinned + 384, blk->offset + rows, sizeof(size_t), cudaMemcpyDeviceToHost, cx->stream));
    WH_CUDA(cudaStreamSynchronize(cx->stream));
    n = *reinterpret_cast<size_t*>(l.pinned + 384);
  }
  WH_REQUIRE(n > 0 && n < (size_t)UINT32_MAX, "wh_worker_begin: %zu entries", n);
  const uint64_t* d_index;
  WH_TRY(to_device(cx, l.in_index, blk->index, n, mem, &d_index));
  WH_TRY(to_device(cx, l.in_offset, blk->offset, rows + 1, mem, &l.b_offset));
  WH_TRY(to_device(cx, l.in_value, blk->value, n, mem, &l.b_value));
  WH_TRY(to_device(cx, l.in_label, blk->label, rows, mem, &l.b_label));
  const size_t ucap = (size_t)std::min<uint64_t>(n, w->max_key);
  WH_TRY(l.uniq.reserve(ucap * 8)); WH_TRY(l.index.reserve(n * 4));
  WH_TRY(l.xw.reserve(rows * 4)); WH_TRY(l.dual.reserve(rows * 4));
  unsigned* d_nuniq = l.scal.as<unsigned>();
  WH_TRY(prof_begin(cx));
  LocInput in{d_index, n, make_xform(w->max_key)};
  if (direct_supported(w->max_key)) WH_TRY(localize_direct(cx, in, l.uniq.as<uint64_t>(), nullptr, l.index.as<uint32_t>(), d_nuniq));
  else WH_TRY(localize_sort(cx, in, l.uniq.as<uint64_t>(), nullptr, l.index.as<uint32_t>(), d_nuniq));
  WH_CUDA(cudaMemcpyAsync(l.pinned + 384, d_nuniq, sizeof(unsigned), cudaMemcpyDeviceToHost, cx->stream));
  WH_CUDA(cudaStreamSynchronize(cx->stream));
  l.rows = rows; l.b_nnz = n; l.b_nuniq = *reinterpret_cast<unsigned*>(l.pinned + 384);
  l.b_open = true;
  *d_uniq = l.uniq.as<uint64_t>();
  *num_uniq = l.b_nuniq;
  return WH_OK;
}

int wh_worker_finish(whWorker* w, const float* d_weights, float* d_grads, int train, double* prog) {
  WH_REQUIRE(w && prog, "wh_worker_finish: NULL argument");
  Lane& l = w->lane[0];
  WH_REQUIRE(l.b_open, "wh_worker_finish: no minibatch begun");
  WH_REQUIRE(l.b_nuniq == 0 || d_weights, "wh_worker_finish: weights are NULL");
  WH_REQUIRE(!train || l.b_nuniq == 0 || d_grads, "wh_worker_finish: training needs a gradient buffer");
  whContext* ctx = w->ctx;
  DeviceGuard g(ctx->device);
  whContext* cx = &l.cx;
  if (!l.cx.own_stream) l.cx.stream = w->ctx->stream;   // lane 0 rides the context's stream, whichever that is now (wh_set_stream)
  l.b_open = false;
  const size_t rows = l.rows, n = l.b_nnz, U = l.b_nuniq;
  FixedScale* fs = reinterpret_cast<FixedScale*>(l.scal.as<char>() + 256);
  EvalOut* d_eval = reinterpret_cast<EvalOut*>(l.scal.as<char>() + 512);
  WH_PROF(cx, "pull_exchange");
  WH_CUDA(cudaMemsetAsync(l.xw.p, 0, rows * 4, cx->stream));
  WH_TRY(spmv_times_launch(cx, rows, l.b_offset, l.index.as<uint32_t>(), l.b_value, d_weights, l.xw.as<float>()));
  WH_PROF(cx, "spmv_times");
  WH_TRY(loss_forward_launch(cx, w->loss, rows, l.b_label, l.xw.as<float>(), train ? l.dual.as<float>() : nullptr, d_eval, 0));
  WH_PROF(cx, "loss_forward");
  WH_CUDA(cudaEventRecord(l.ev_fork, cx->stream));
  WH_CUDA(cudaStreamWaitEvent(l.ax.stream, l.ev_fork, 0));
  WH_TRY(auc_launch(&l.ax, rows, l.b_label, l.xw.as<float>(), d_eval));
  WH_CUDA(cudaEventRecord(l.ev_auc, l.ax.stream));
  if (train && U) {
    long long* acc;
    WH_TRY(ensure_acc(cx, U, &acc));
    WH_TRY(spmv_trans_accumulate_launch(cx, rows, n, l.b_offset, l.index.as<uint32_t>(), l.b_value, l.dual.as<float>(), acc, fs, acc_copies(), U));
    WH_PROF(cx, "spmv_trans");
    WH_TRY(fixed_finish_launch(cx, acc, U, fs, d_grads, w->loss == WH_LOSS_SQUARE_HINGE ? -2.f : 1.f, acc_copies(), U));
    WH_PROF(cx, "grad_finish");
  }
  WH_CUDA(cudaStreamWaitEvent(cx->stream, l.ev_auc, 0));
  if (cx->prof.on) WH_TRY(prof_mark(cx, "auc_tail"));
  WH_CUDA(cudaMemcpyAsync(l.pinned, d_eval, sizeof(EvalOut), cudaMemcpyDeviceToHost, cx->stream));
  WH_CUDA(cudaStreamSynchronize(cx->stream));
  if (cx->prof.on) {
    WH_TRY(prof_collect(cx));
    for (auto& a : cx->prof.acc) {
      Profiler::Acc* t = nullptr;
      for (auto& x : ctx->prof.acc) if (x.name == a.name) { t = &x; break; }
      if (!t) { ctx->prof.acc.push_back(Profiler::Acc()); t = &ctx->prof.acc.back(); t->name = a.name; }
      t->ms += a.ms; t->count += a.count;
    }
    cx->prof.acc.clear();
  }
  EvalHost h = eval_finish(*reinterpret_cast<const EvalOut*>(l.pinned), rows);
  prog[0] = h.objv; prog[1] = h.acc; prog[2] = h.auc; prog[3] = 1; prog[4] = (double)rows; prog[5] = 0;
  return WH_OK;
}

int wh_worker_process_minibatch(whWorker* w, const whRowBlock* blk, int mem, int train, double* prog) {
  WH_REQUIRE(w && blk && prog, "wh_worker_process_minibatch: NULL argument");
  for (int i = 0; i < 6; ++i) prog[i] = 0;
  if (blk->size == 0) return WH_OK;
  WH_REQUIRE(w->next_wait == w->next_ticket, "wh_worker_process_minibatch: minibatches submitted earlier are still in flight");
  WH_REQUIRE(w->server != nullptr, "wh_worker_process_minibatch: this worker has no local server; use wh_worker_begin / wh_worker_finish");
  uint64_t t = 0;
  WH_TRY(wh_worker_submit(w, blk, mem, 0, train, &t));
  return wh_worker_wait(w, t, prog);
}

}  // extern "C"
