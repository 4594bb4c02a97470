// The per-minibatch pipeline of one worker stream: AsgdWorker::ProcessMinibatch
// (learn/linear/async_sgd.h:240-288) with the server on the same GPU.
//
//   Localize (:246-247) -> ZPull (:287) -> Init = X w (:260) -> Evaluate (:261) ->
//   CalcGrad = X^T dual (:269) -> ZPush (:277-278)
//
// Everything is enqueued on the context's stream without host round trips: the number of unique
// keys U stays on the device (kernels read it there), and the only synchronisation is the final
// read-back of the progress numbers.
#include "batch.cuh"
#include "loss.cuh"
#include "server.cuh"
#include "spmv.cuh"

struct whWorker {
  whContext* ctx = nullptr;
  whServer* server = nullptr;
  int loss = WH_LOSS_LOGIT;
  uint64_t max_key = WH_NO_MAX_KEY;
  long long seen_new_w = 0;
  wh::DevBuf in_index, in_offset, in_value, in_label;   // staged host input
  wh::DevBuf uniq, index, w, xw, dual, scal;
};

namespace wh {
int spmv_trans_accumulate_launch(whContext* ctx, size_t rows, size_t nnz, const size_t* offset,
                                 const uint32_t* index, const float* value, const float* x,
                                 long long* acc, FixedScale* fs);
}

using namespace wh;

extern "C" {

int wh_worker_create(whContext* ctx, whServer* server, int loss, uint64_t max_key, whWorker** out) {
  WH_REQUIRE(ctx && server && out, "wh_worker_create: NULL argument");
  WH_REQUIRE(server->ctx == ctx, "wh_worker_create: server lives on another context");
  WH_REQUIRE(loss == WH_LOSS_LOGIT || loss == WH_LOSS_SQUARE_HINGE,
             "wh_worker_create: unknown loss type %d (SQUARE is unimplemented in the reference, learn/linear/loss.h:163-180)", loss);
  WH_REQUIRE(max_key != 0, "wh_worker_create: max_key must be positive");
  whWorker* w = new whWorker();
  w->ctx = ctx; w->server = server; w->loss = loss; w->max_key = max_key;
  DeviceGuard g(ctx->device);
  int r = w->scal.reserve(4096);
  if (r != WH_OK) { delete w; return r; }
  *out = w;
  return WH_OK;
}

int wh_worker_destroy(whWorker* w) {
  if (!w) return WH_OK;
  DeviceGuard g(w->ctx->device);
  cudaStreamSynchronize(w->ctx->stream);
  for (wh::DevBuf* b : {&w->in_index, &w->in_offset, &w->in_value, &w->in_label, &w->uniq, &w->index, &w->w,
                        &w->xw, &w->dual, &w->scal}) b->release();
  delete w;
  return WH_OK;
}

int wh_worker_process_minibatch(whWorker* w, const whRowBlock* blk, int mem, int train, double* prog) {
  WH_REQUIRE(w && blk && prog, "wh_worker_process_minibatch: NULL argument");
  WH_REQUIRE(mem == WH_HOST || mem == WH_DEVICE, "wh_worker_process_minibatch: bad mem %d", mem);
  whContext* ctx = w->ctx;
  whServer* s = w->server;
  DeviceGuard g(ctx->device);
  for (int i = 0; i < 6; ++i) prog[i] = 0;
  const size_t rows = blk->size;
  if (rows == 0) return WH_OK;
  WH_REQUIRE(blk->offset && blk->index && blk->label, "wh_worker_process_minibatch: offset/index/label is NULL");
  size_t n;
  if (mem == WH_HOST) n = blk->offset[rows];
  else {
    WH_CUDA(cudaMemcpyAsync(ctx->pinned.p, blk->offset + rows, sizeof(size_t), cudaMemcpyDeviceToHost, ctx->stream));
    WH_CUDA(cudaStreamSynchronize(ctx->stream));
    n = ctx->pinned.as<size_t>()[0];
  }
  WH_REQUIRE(n > 0 && n < (size_t)UINT32_MAX, "wh_worker_process_minibatch: %zu entries", n);
  const bool direct = direct_supported(w->max_key);

  // stage the minibatch
  const uint64_t* d_index; const size_t* d_offset; const float* d_value; const float* d_label;
  WH_TRY(to_device(ctx, w->in_index, blk->index, n, mem, &d_index));
  WH_TRY(to_device(ctx, w->in_offset, blk->offset, rows + 1, mem, &d_offset));
  WH_TRY(to_device(ctx, w->in_value, blk->value, n, mem, &d_value));
  WH_TRY(to_device(ctx, w->in_label, blk->label, rows, mem, &d_label));

  const size_t ucap = (size_t)std::min<uint64_t>(n, w->max_key);
  WH_TRY(w->uniq.reserve(ucap * 8)); WH_TRY(w->index.reserve(n * 4)); WH_TRY(w->w.reserve(ucap * 4));
  WH_TRY(w->xw.reserve(rows * 4)); WH_TRY(w->dual.reserve(rows * 4));
  unsigned* d_nuniq = w->scal.as<unsigned>();
  FixedScale* fs = reinterpret_cast<FixedScale*>(w->scal.as<char>() + 256);
  EvalOut* d_eval = reinterpret_cast<EvalOut*>(w->scal.as<char>() + 512);

  WH_TRY(prof_begin(ctx));
  // Localize
  LocInput in{d_index, n, make_xform(w->max_key)};
  if (direct) WH_TRY(localize_direct(ctx, in, w->uniq.as<uint64_t>(), nullptr, w->index.as<uint32_t>(), d_nuniq));
  else WH_TRY(localize_sort(ctx, in, w->uniq.as<uint64_t>(), nullptr, w->index.as<uint32_t>(), d_nuniq));
  // ZPull: one Handle::Pull per unique key
  WH_TRY(server_pull_launch(s, w->uniq.as<uint64_t>(), ucap, d_nuniq, w->w.as<float>()));
  WH_PROF(ctx, "pull");
  // Init: Xw
  WH_CUDA(cudaMemsetAsync(w->xw.p, 0, rows * 4, ctx->stream));
  WH_TRY(spmv_times_launch(ctx, rows, d_offset, w->index.as<uint32_t>(), d_value, w->w.as<float>(), w->xw.as<float>()));
  WH_PROF(ctx, "spmv_times");
  // Evaluate (+ the dual of CalcGrad, which falls out of the same pass)
  WH_TRY(loss_forward_launch(ctx, w->loss, rows, d_label, w->xw.as<float>(), train ? w->dual.as<float>() : nullptr, d_eval, 0));
  WH_PROF(ctx, "loss_forward");
  WH_TRY(auc_launch(ctx, rows, d_label, w->xw.as<float>(), d_eval));
  WH_PROF(ctx, "auc");
  if (train) {
    // CalcGrad: X^T dual into the fixed-point accumulator; ZPush: convert + update in one pass
    long long* acc;
    WH_TRY(ensure_acc(ctx, ucap, &acc));
    WH_TRY(spmv_trans_accumulate_launch(ctx, rows, n, d_offset, w->index.as<uint32_t>(), d_value, w->dual.as<float>(), acc, fs));
    WH_PROF(ctx, "spmv_trans");
    WH_TRY(server_push_acc_launch(s, w->uniq.as<uint64_t>(), ucap, d_nuniq, acc, fs, w->loss == WH_LOSS_SQUARE_HINGE ? -2.f : 1.f));
  }
  if (train) WH_PROF(ctx, "push");
  // progress read-back (the one synchronisation of the step)
  char* pin = ctx->pinned.as<char>();
  WH_CUDA(cudaMemcpyAsync(pin, d_eval, sizeof(EvalOut), cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaMemcpyAsync(pin + 256, s->counters.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->prof.on) WH_TRY(prof_collect(ctx));
  const long long* cnt = reinterpret_cast<const long long*>(pin + 256);
  if (cnt[1]) {
    WH_CUDA(cudaMemsetAsync(s->counters.as<long long>() + 1, 0, 8, ctx->stream));
    set_error("wh_worker_process_minibatch: %lld keys fall outside the server's range [%llu, %llu)", cnt[1],
              (unsigned long long)s->key_begin, (unsigned long long)s->key_end);
    return WH_ERR_INVALID;
  }
  EvalHost h = eval_finish(*reinterpret_cast<const EvalOut*>(pin), rows);
  prog[0] = h.objv; prog[1] = h.acc; prog[2] = h.auc; prog[3] = 1; prog[4] = (double)rows;
  prog[5] = (double)(cnt[0] - w->seen_new_w);
  w->seen_new_w = cnt[0];
  return WH_OK;
}

}  // extern "C"
