/*
 * wormhole_b200.h -- C ABI of libwormhole_b200.so
 *
 * The sparse minibatch training hot path of dmlc wormhole (learn/linear async SGD /
 * AdaGrad / FTRL, shared by learn/difacto) as hand-written sm_100a CUDA kernels.
 * The reference has no FFI for this path: its seams are C++ header interfaces
 * compiled into each app.  Every entry point below names the reference interface
 * it stands in for (paths relative to the reference tree).  The header-only C++
 * facade in wormhole_b200/host/ gives them back the reference's class names and
 * signatures; INTEGRATION.md shows the binding a maintainer would add.
 *
 * Conventions
 *   - plain pointers and sizes only; no CUDA or torch types in signatures
 *     (a CUDA stream crosses as void*).
 *   - `mem` says where EVERY array argument of that call lives: WH_HOST (pageable
 *     or pinned host memory; the call copies in, computes on the GPU, copies out,
 *     and is synchronous) or WH_DEVICE (device pointers on the context's device;
 *     the call only enqueues work on the context's stream -- wh_sync() to wait).
 *   - every function returns WH_OK (0) or an error code; wh_last_error() gives the
 *     message of the calling thread's last failure.  There is no CPU fallback: a
 *     call that cannot run on the GPU fails.
 *   - one whContext per GPU; calls on one context must be serialised by the caller.
 */
#ifndef WORMHOLE_B200_H_
#define WORMHOLE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WH_API __attribute__((visibility("default")))

enum { WH_OK = 0, WH_ERR_INVALID = 1, WH_ERR_CUDA = 2, WH_ERR_NOMEM = 3, WH_ERR_UNSUPPORTED = 4 };
enum { WH_HOST = 0, WH_DEVICE = 1 };
/* learn/linear/config.proto:30-38 (Config::Loss) */
enum { WH_LOSS_SQUARE = 1, WH_LOSS_LOGIT = 2, WH_LOSS_SQUARE_HINGE = 4 };
/* learn/linear/config.proto:45-52 (Config::Algo) */
enum { WH_ALGO_SGD = 1, WH_ALGO_ADAGRAD = 2, WH_ALGO_FTRL = 3 };
/* localizer strategy */
enum { WH_LOC_AUTO = 0, WH_LOC_DIRECT = 1, WH_LOC_SORT = 2,
       WH_LOC_WITH_COUNTS = 16 /* OR into method: also produce the per-key counts (idx_frq) */ };
/* ps::FLAGS_max_key default: no hashing, keys are nibble-reversed ids (learn/base/localizer.h:10,108-121) */
#define WH_NO_MAX_KEY UINT64_MAX

typedef struct whContext whContext; /* one GPU: device, stream, scratch */
typedef struct whBatch whBatch;     /* a localized minibatch, device resident */
typedef struct whServer whServer;   /* one key range of the model, HBM resident */
typedef struct whWorker whWorker;   /* the per-minibatch pipeline of one worker stream */

/* dmlc::RowBlock<uint64_t> by value: what MinibatchIter hands to ProcessMinibatch
 * (learn/solver/minibatch_solver.h:211, learn/base/minibatch_iter.h:123-125).
 * offset[0] must be 0 (learn/base/localizer.h:102); value may be NULL (all ones). */
typedef struct {
  size_t size;           /* rows */
  const size_t* offset;  /* size+1 */
  const float* label;    /* size, may be NULL */
  const float* weight;   /* unused by this path */
  const uint64_t* index; /* offset[size] feature ids */
  const float* value;    /* offset[size] or NULL */
} whRowBlock;

/* dmlc::RowBlock<unsigned>: a localized block (learn/base/spmv.h:48, SpMV::SpMat). */
typedef struct {
  size_t size;
  const size_t* offset;
  const float* label;
  const uint32_t* index;
  const float* value;
} whLocalBlock;

/* ---- context -------------------------------------------------------------------------- */
WH_API const char* wh_last_error(void);
WH_API const char* wh_version(void);
/* Number of kernels this library has launched from the calling process (all contexts). */
WH_API uint64_t wh_kernel_launches(void);
WH_API int wh_create(int device, whContext** out);
WH_API int wh_destroy(whContext* ctx);
/* Use an existing CUDA stream (cudaStream_t as void*) instead of the context's own. Synchronises the old stream first.
 * Workers created on the context follow: their first lane rides the context's stream, whichever it is at each call.
 * Callers that mix their own device work with WH_DEVICE calls (which only enqueue) must use this stream for it. */
WH_API int wh_set_stream(whContext* ctx, void* cuda_stream);
WH_API int wh_sync(whContext* ctx);
/* Opt-in timing of the library's own kernels with CUDA events on the context's stream (what
 * bench.py's roofline figures are measured with). Enabling clears the accumulated numbers.
 * Segment i: name, summed milliseconds, number of times it ran. No counterpart in the reference
 * beyond its wall-clock workload_time_ (learn/linear/async_sgd.h:245-248,258,285). */
WH_API int wh_profile_enable(whContext* ctx, int on);
WH_API int wh_profile_count(whContext* ctx);
WH_API int wh_profile_get(whContext* ctx, int i, const char** name, double* total_ms, uint64_t* count);

/* ---- (1) localizer ---------------------------------------------------------------------- */
/* dmlc::ReverseBytes, learn/base/localizer.h:16-26 (a nibble reversal). Host-side scalar helper
 * for callers that need to name a key (model files, key ranges); the kernels apply it on device. */
WH_API uint64_t wh_reverse_bytes(uint64_t x);

/* Localizer<uint64_t>::Localize, learn/base/localizer.h:42-51 (= CountUniqIndex :96-167 then
 * RemapIndex :169-221 against the block's own unique list). max_key is ps::FLAGS_max_key.
 * method: WH_LOC_AUTO picks direct addressing for hashed key spaces and the radix sort otherwise.
 * The result stays on the device; read it with wh_batch_*. */
WH_API int wh_localize(whContext* ctx, const whRowBlock* blk, int mem, uint64_t max_key, int method,
                       whBatch** out);
/* CountUniqIndex(blk) then RemapIndex(blk, dict): entries whose transformed key is not in the
 * ascending dictionary are dropped and the CSR is compacted (learn/base/localizer.h:179-213).
 * wh_batch uniq/cnt then describe blk itself, index/offset/value the remapped block. */
WH_API int wh_localize_with_dict(whContext* ctx, const whRowBlock* blk, int mem, uint64_t max_key,
                                 const uint64_t* dict, size_t ndict, whBatch** out);
WH_API size_t wh_batch_rows(const whBatch* b);
WH_API size_t wh_batch_nnz(const whBatch* b);
WH_API size_t wh_batch_num_uniq(const whBatch* b);
WH_API int wh_batch_has_value(const whBatch* b);
WH_API int wh_batch_has_label(const whBatch* b);
/* Copies the parts asked for (NULL = skip) to host arrays sized by the getters above:
 * uniq[U] ascending transformed keys, cnt_u32[U] multiplicities (C = unsigned), cnt_f32[U] the
 * same cast to float (C = float, learn/difacto/async_sgd.h:366), offset[rows+1], index[nnz],
 * value[nnz], label[rows]. */
WH_API int wh_batch_copy(whContext* ctx, const whBatch* b, uint64_t* uniq, uint32_t* cnt_u32,
                         float* cnt_f32, size_t* offset, uint32_t* index, float* value, float* label);
/* Device views of the same arrays (valid until wh_batch_destroy). */
WH_API int wh_batch_device_view(const whBatch* b, const uint64_t** uniq, const uint32_t** cnt,
                                whLocalBlock* local);
WH_API int wh_batch_destroy(whBatch* b);

/* ---- (2) SpMV, (5) SpMM ----------------------------------------------------------------- */
/* SpMV::Times, learn/base/spmv.h:72-91: y[i] = sum_j x[index[j]] * value[j] in CSR order
 * (sequential fp32, bit-exact with the reference); rows without entries leave y[i] untouched. */
WH_API int wh_spmv_times(whContext* ctx, const whLocalBlock* D, int mem, const float* x, size_t xsize,
                         float* y);
/* SpMV::TransTimes, learn/base/spmv.h:95-119: y[0..ysize) = D^T x. */
WH_API int wh_spmv_trans_times(whContext* ctx, const whLocalBlock* D, int mem, const float* x,
                               float* y, size_t ysize);
/* Allocation-free variant of wh_spmv_trans_times. It has no counterpart in the reference.
 * For callers that must not allocate device memory: a worker whose GPU also runs other minibatch
 * streams (cudaMalloc / cudaFree stall the whole device) or a process whose device memory is
 * already fully owned by someone else (cudaMalloc fails). Needs no whContext and calls no
 * cudaMalloc, cudaFree or any allocating routine: its scratch (the fixed-point accumulators) is a
 * static 2 MiB array inside the library, held by every GPU that loads it until the process exits.
 *   - all arrays are DEVICE pointers on the current device; work is enqueued on `cuda_stream`
 *     (cudaStream_t as void*, NULL = default stream); nnz = offset[D->size], passed by the caller
 *     because reading it back would need a synchronising copy.
 *   - ysize must not exceed WH_STATIC_SCRATCH_MAX_OUTPUTS; larger requests fail with
 *     WH_ERR_INVALID and a message naming the limit.
 *   - every D->index[j] must be < ysize (true of any block localized against ysize keys). The call
 *     cannot check this without reading the indices back; an index beyond ysize would add into the
 *     library's static array past the cleared range, or past the array itself.
 *   - the first call on each device looks up the array's device address; make that first call
 *     before any stream capture begins.
 *   - the array is shared by every call on that GPU and keeps the last call's contents (each call
 *     clears the part it uses first). Calls that overlap on the same GPU race: serialise them. */
#define WH_STATIC_SCRATCH_MAX_OUTPUTS 262144
WH_API int wh_spmv_trans_times_static(void* cuda_stream, const whLocalBlock* D, size_t nnz, const float* x,
                                      float* y, size_t ysize);
/* SpMM::Times, learn/base/spmm.h:23-29,56-81: Y[rows x dim] = D X[xrows x dim], row major. */
WH_API int wh_spmm_times(whContext* ctx, const whLocalBlock* D, int mem, const float* x, size_t xsize,
                         float* y, int dim);
/* SpMM::TransTimes, learn/base/spmm.h:34-52,85-123: Y[ysize] = p*Z + D^T X (z may be NULL). */
WH_API int wh_spmm_trans_times(whContext* ctx, const whLocalBlock* D, int mem, const float* x, float p,
                               const float* z, float* y, size_t ysize, int dim);

/* ---- (3) loss --------------------------------------------------------------------------- */
/* BinClassEval<float>, learn/base/binary_class_evaluation.h:17-74.
 * out = {AUC, Accuracy(0), LogLoss, LogitObjv, AUC_lo, AUC_hi}. Examples with equal scores are
 * left in unspecified order by the reference's std::sort (:25-26); AUC counts a tied
 * (positive, negative) pair as 1/2 and AUC_lo / AUC_hi bound every possible tie order. */
WH_API int wh_bin_eval(whContext* ctx, const float* label, const float* predict, size_t n, int mem,
                       float* out6);
/* CreateLoss(loss)->Init(D, w, nt); Evaluate(&prog); CalcGrad(&grad)
 * (learn/linear/loss.h:32-49,81-115,128-156,172-182).
 * prog[6] = {objv, acc, auc, count, new_ex, new_w} (learn/linear/progress.h:28-34), host memory.
 * grad[U] (NULL: skip CalcGrad), xw[rows] and dual[rows] (NULL: not returned) follow `mem`.
 * WH_LOSS_SQUARE: declared by the reference (learn/linear/config.proto:30-34, "1/2 (p - y)^2") but never implemented
 * there (loss.h:163-166 is an empty TODO, CreateLoss :178-180 aborts). Implemented here as the declaration reads, with
 * y = +-1 as in the other losses: dual = p - y, objective = sum 1/2 (p - y)^2. It has no reference to be checked against. */
WH_API int wh_linear_loss(whContext* ctx, int loss, const whLocalBlock* D, int mem, const float* w,
                          size_t U, double* prog, float* grad, float* xw, float* dual);

/* ---- (4) server side: the model shard and its update handles ---------------------------- */
typedef struct {
  int algo;                /* WH_ALGO_* : SGDHandle / AdaGradHandle / FTRLHandle */
  float lambda_l1;         /* conf.lambda_l1(), learn/linear/async_sgd.h:203 */
  float lambda_l2;         /* conf.lambda_l2(), :204 */
  int has_lr_eta;          /* handle alpha stays 0.1 unless the key is present (:49,205) */
  float lr_eta;
  int has_lr_beta;         /* handle beta stays 1 unless the key is present (:49,206) */
  float lr_beta;
  uint64_t key_begin;      /* this shard owns transformed keys in [key_begin, key_end) */
  uint64_t key_end;        /* dense direct-addressed table: key_end - key_begin entries */
  uint64_t hash_slots;     /* 0: dense. Else an open-addressing table of this many slots (power of two) for
                              un-hashed 64-bit keys -- ps::FLAGS_max_key at its default, keys = nibble-reversed
                              ids spread over the whole 64-bit range; key_begin / key_end still bound the range */
} whServerConfig;

/* AsgdServer::CreateServer<Entry,Handle>, learn/linear/async_sgd.h:200-213. */
WH_API int wh_server_create(whContext* ctx, const whServerConfig* conf, whServer** out);
WH_API int wh_server_destroy(whServer* s);
/* One pull message: Handle::Pull per key (learn/linear/async_sgd.h:97-99,137-139,177-179). */
WH_API int wh_server_pull(whServer* s, const uint64_t* keys, size_t n, int mem, float* vals);
/* One push message: Handle::Start(push) (SGD: eta from t, t += 1, :85-90), Handle::Push per key
 * in message order (:91-95,123-135,160-175), Handle::Finish. Keys must be distinct (they are the
 * unique list of one minibatch). */
WH_API int wh_server_push(whServer* s, const uint64_t* keys, size_t n, int mem, const float* grads);
/* ISGDHandle::new_w accumulated so far: the number of non-zero weights (:35-41). Synchronises. */
WH_API int wh_server_nnz_w(whServer* s, int64_t* out);
/* Entry fields for the given keys: w, z (FTRL), sq_cum_grad (AdaGrad/FTRL); NULL = skip. */
WH_API int wh_server_get_state(whServer* s, const uint64_t* keys, size_t n, int mem, float* w, float* z,
                               float* sq_cum_grad);
WH_API int wh_server_set_state(whServer* s, const uint64_t* keys, size_t n, int mem, const float* w,
                               const float* z, const float* sq_cum_grad);
/* AsgdServer::SaveModel / LoadModel (learn/linear/async_sgd.h:215-223 -> KVStore::Save/Load -> Entry::Save/Load,
 * :59-66): one {uint64 key, float w} record per entry with w != 0, ascending keys (ps-lite's own record framing is
 * not in the reference tree). The caller names the file as IterServer::ModelName does
 * (<model_out>[_iter-k]_part-<rank>, learn/solver/iter_solver.h:115-119). Load restores w only and counts it in
 * |w|_0, as TLoad does; optimizer state is not checkpointed by the reference either. */
WH_API int wh_server_save(whServer* s, const char* path);
WH_API int wh_server_load(whServer* s, const char* path);
/* SGDHandle::t (learn/linear/async_sgd.h:100). */
WH_API int wh_server_sgd_t(const whServer* s);
/* Key-range shards on several GPUs of one box (one process per GPU). ps-lite gives server g an even slice of
 * [0, max_key) and routes each worker's key slices to it over ZeroMQ; here every rank maps every other rank's
 * shard over NVLink (CUDA IPC) and the pull / push kernels of a worker address the owning GPU directly.
 * export: fills WH_IPC_HANDLE_BYTES bytes that the caller passes to the other ranks by any means.
 * attach: `handles` = the exports of ranks 0..world-1 concatenated; this shard must have been created with the
 * range [rank*max_key/world, (rank+1)*max_key/world). Afterwards a worker created on this server runs the sharded
 * ProcessMinibatch: gather from the owners, barrier, then `world` push turns (turn k updates shard (rank+k)%world)
 * with stream-ordered barriers in between; every rank must process the same number of minibatches. */
#define WH_IPC_HANDLE_BYTES 192
WH_API int wh_server_ipc_export(whServer* s, void* handle);
WH_API int wh_server_ipc_attach(whServer* s, int rank, int world, uint64_t max_key, const void* handles);

/* ---- the per-minibatch pipeline --------------------------------------------------------- */
/* AsgdWorker, learn/linear/async_sgd.h:229-305, bound to one server shard that owns every key
 * (1 worker, 1 server). max_key is ps::FLAGS_max_key. */
WH_API int wh_worker_create(whContext* ctx, whServer* server, int loss, uint64_t max_key,
                            whWorker** out);
/* Same, with `depth` minibatches allowed in flight (conf max_concurrency,
 * learn/solver/minibatch_solver.h:289-290,307-322); depth in [1, 8]. */
WH_API int wh_worker_create_ex(whContext* ctx, whServer* server, int loss, uint64_t max_key, int depth,
                               whWorker** out);
WH_API int wh_worker_destroy(whWorker* w);
/* The two halves of ProcessMinibatch for callers that keep several minibatches in flight
 * (MinibatchWorker::Process / WaitMinibatch / FinishMinibatch, learn/solver/minibatch_solver.h:
 * 261-276,307-322). submit enqueues the whole minibatch and returns at once; nnz = offset[size]
 * (only read for WH_DEVICE blocks; 0 = unknown, costs a synchronising read); the arrays of blk must
 * stay valid and unchanged until the matching wait. wait blocks until that minibatch is done and
 * fills prog[6]. Tickets are collected in submission order. Every pull sees all earlier pushes, so
 * the numbers equal strictly sequential processing; staging, localize and AUC of neighbouring
 * minibatches overlap. (Binary features in a hashed key space take the worker's hot-key path -- a cache of the
 * stream's most frequent keys, rebuilt from time to time; it changes the speed, never the numbers.) */
/* Bounded staleness. 0 (default): every pull sees all earlier pushes -- the numbers of sequential processing,
 * the configuration parity is judged in ("max_delay = 0"). d > 0: the pull of minibatch t only waits for the push
 * of minibatch t-1-d, as in the reference when max_concurrency minibatches are in flight and a pull overtakes the
 * worker's own earlier pushes (learn/solver/minibatch_solver.h:307-311); pushes are still applied in order.
 * Needs d < depth. Sharded jobs: after wh_worker_ipc_attach only, and the same d on every rank. */
WH_API int wh_worker_set_max_delay(whWorker* w, int max_delay);
/* The worker's table of frequent keys (a cache of the stream's most popular transformed keys, built from a census of one
 * of its minibatches; it changes the speed, never the numbers): the keys it holds (up to cap; *n = how many there are),
 * the wall-clock cost of the last census and how many were taken. No counterpart in the reference; bench.py reports the
 * hit rate with it. */
WH_API int wh_worker_table_keys(whWorker* w, uint32_t* keys, size_t cap, size_t* n, double* census_ms, int* censuses);
WH_API int wh_worker_submit(whWorker* w, const whRowBlock* blk, int mem, size_t nnz, int train,
                            uint64_t* ticket);
WH_API int wh_worker_wait(whWorker* w, uint64_t ticket, double* prog);
/* ProcessMinibatch split at its ZPull (learn/linear/async_sgd.h:246-247 | :257-286), for servers that
 * live elsewhere (key-range shards on other GPUs; the worker may be created with server = NULL).
 * begin: Localize; returns a DEVICE pointer to the ascending unique keys (valid until the next begin)
 * and their number -- the key list of the pull and the push. Synchronises once to learn the number.
 * finish: the pull callback -- Init / Evaluate / CalcGrad with the pulled weights d_weights[num_uniq]
 * (device); the gradient is written to d_grads[num_uniq] (device) for the caller's push. prog[5]
 * (new_w) is 0: the servers own that count (wh_server_nnz_w). */
/* Sharded jobs, owner-side push. After wh_server_ipc_attach, a worker can publish one outbox per lane ({key
 * list, gradients, inbox of pulled weights}; capacity max_uniq keys, 64 bytes of handle per lane written to `handles`) and map its peers'
 * (`handles` = all ranks' exports in rank order, world x depth x 64 bytes; same depth everywhere). Pull and push are
 * then served by the OWNERS: once every rank has posted its key list, each owner writes the weights of the slices it
 * owns into the requesters' inboxes; once every rank has posted its gradients, each owner applies, per source rank in
 * ascending order, the slice of that rank's outbox that falls into its key range as one push message. Every rank must
 * submit the same sequence of minibatches (train / evaluate alike): the barriers in between are collective. */
/* Ragged streams: the ranks of a sharded job read different file parts and run out of minibatches at different steps,
 * while the barriers between the GPUs are collective. A rank WITHOUT a minibatch in a step calls submit_idle instead of
 * submit (same train flag as the others): it posts an empty key list, serves the others' pulls and applies their pushes;
 * its ticket's progress is all zero. Before each step every rank tells its server which ranks do send a push message
 * (bit r = rank r has a minibatch), so that SGD's per-message step count t advances for those only
 * (learn/linear/async_sgd.h:85-90). The host agrees on the mask by its own means (tools/linear_b200.cc: one pipe
 * round-trip per step). Default: all ranks. */
WH_API int wh_worker_submit_idle(whWorker* w, int train, uint64_t* ticket);
WH_API int wh_server_set_active_sources(whServer* s, uint64_t mask);
WH_API int wh_worker_ipc_export(whWorker* w, size_t max_uniq, void* handles);
WH_API int wh_worker_ipc_attach(whWorker* w, const void* handles);
/* The PRED workload (learn/linear/async_sgd.h:263-265 -> ScalarLoss::Predict, learn/linear/loss.h:55-63): Localize,
 * ZPull, Xw; out[rows] (host) receives 1/(1+exp(-Xw)) if prob_predict else Xw. Needs a local, unsharded server. */
WH_API int wh_worker_predict(whWorker* w, const whRowBlock* blk, int mem, int prob_predict, float* out);
WH_API int wh_worker_begin(whWorker* w, const whRowBlock* blk, int mem, size_t nnz, const uint64_t** d_uniq,
                           size_t* num_uniq);
WH_API int wh_worker_finish(whWorker* w, const float* d_weights, float* d_grads, int train, double* prog);
/* AsgdWorker::ProcessMinibatch, learn/linear/async_sgd.h:240-288: Localize -> ZPull ->
 * Init/Evaluate -> (train) CalcGrad -> ZPush. prog[6] as in wh_linear_loss, with new_w the
 * server's change of |w|_0 for this minibatch. Synchronous. */
WH_API int wh_worker_process_minibatch(whWorker* w, const whRowBlock* blk, int mem, int train,
                                       double* prog);

/* ---- (4b) the reader in front of the path ------------------------------- */
/* Formats of MinibatchIter (learn/base/minibatch_iter.h:44-59): the text formats below through wh_parse_text, "crb"
 * (CompressedRowBlock records in RecordIO framing) through wh_parse_crb. */
enum { WH_FMT_LIBSVM = 0, WH_FMT_CRITEO = 1, WH_FMT_CRITEO_TEST = 2, WH_FMT_ADFEA = 3 };
typedef struct whRows whRows; /* a device-resident dmlc::data::RowBlockContainer<uint64_t> */
/* ParseNext of one chunk of text (whole lines; host or device memory, < 4 GiB) on the GPU:
 *   criteo / criteo_test  CriteoParser::ParseNext, learn/base/criteo_parser.h:38-89: label, 13 integer fields and 26
 *                         8-byte categorical fields, tab separated; index = CityHash64(field) >> 10 | field_no << 54.
 *   libsvm                dmlc-core's LibSVMParser (un-vendored): `label index[:value] ...`, numbers as atof / strtoull.
 *   adfea                 AdfeaParser::ParseNext, learn/base/adfea_parser.h:38-82, one row per line: `lineid count label
 *                         idx:gid ...`; index = idx >> 10 | gid << 54, label = (first character of the label token == '1').
 * Where the reference would read past the end of a line or CHECK-fail (short lines, categorical fields that are not
 * 8 bytes) the call fails with WH_ERR_INVALID and names the line. Numeric literals are converted like strtod / atof
 * (correctly rounded; literals of 16 and more digits or large exponents through exact big-integer arithmetic); only
 * inf / nan / hexadecimal literals are refused.
 * *out == NULL on entry: a new container is returned; otherwise that container is cleared and refilled, reusing
 * its device buffers (no allocation in steady state). */
WH_API int wh_parse_text(whContext* ctx, int format, const char* text, size_t len, int mem, whRows** out);
/* CRBParser::ParseNext, learn/base/crb_parser.h:27-40, for every record of a chunk: `data` (HOST memory) holds whole
 * RecordIO records (dmlc/recordio.h framing: magic 0xced7230a, {cflag:3 | length:29}, payload padded to 4 bytes; split
 * records are re-joined), each a CompressedRowBlock (learn/base/compressed_row_block.h:44-69,116-137: magic, index
 * size -- 8 here --, row count, then label / offset / index / value / weight as {int size, LZ4 block}). The LZ4 blocks
 * are decoded on the host (a strictly sequential byte code) and the rows appended to the device container; a value
 * array of all ones was already dropped by the writer (:27-35). Same *out convention as wh_parse_text. */
WH_API int wh_parse_crb(whContext* ctx, const char* data, size_t len, whRows** out);
/* The LZ4 block decoder wh_parse_crb uses in place of lz4's LZ4_decompress_safe (compressed_row_block.h:121-123; lz4 is
 * not vendored by the reference): decodes one block into dst[0, cap), returns the decoded size or -1 on malformed
 * input. Host memory, no GPU involved. */
WH_API long wh_lz4_block_decode(const void* src, size_t n, void* dst, size_t cap);
/* RowBlockContainer: Clear / Push(RowBlock slice) / Push(Row) / GetBlock, as MinibatchIter::Next uses them
 * (learn/base/minibatch_iter.h:74-147). append takes rows [begin, begin+n) of src, or the n rows listed in the HOST
 * array row_ids (the shuffled order rdp_[i], :102-110). drop_unit_values is the all-ones rule (:116-118).
 * view yields DEVICE pointers, ready for wh_worker_submit(..., WH_DEVICE, wh_rows_nnz(r), ...). */
WH_API int wh_rows_create(whContext* ctx, whRows** out);
WH_API void wh_rows_destroy(whRows* r);
WH_API int wh_rows_clear(whRows* r);
WH_API int wh_rows_append(whContext* ctx, whRows* dst, const whRows* src, size_t begin, size_t n, const uint32_t* row_ids);
WH_API int wh_rows_drop_unit_values(whContext* ctx, whRows* r);
WH_API size_t wh_rows_size(const whRows* r);
WH_API size_t wh_rows_nnz(const whRows* r);
WH_API int wh_rows_has_value(const whRows* r);
WH_API int wh_rows_view(const whRows* r, whRowBlock* view);
/* to host arrays (NULL = skip): offset[size+1], label[size], index[nnz], value[nnz] */
WH_API int wh_rows_copy(whContext* ctx, const whRows* r, size_t* offset, float* label, uint64_t* index, float* value);

/* ---- (5) DiFacto: factorization machine on the same path -------------------------------- */
typedef struct whDifacto whDifacto; /* AsyncServer shard + AsyncWorker pipeline on one GPU */

/* The conf keys AsyncServer / AsyncWorker / Loss read (learn/difacto/config.proto:6-167,
 * learn/difacto/async_sgd.h:302-327,344-357, learn/difacto/loss.h:28-40). */
typedef struct {
  float lambda_l1, lambda_l2;   /* w */
  float lr_eta, lr_beta;        /* w: alpha, beta of the FTRL-style UpdateW */
  int l1_shrk;                  /* conf.l1_shrk(): pull / grow embeddings only where w != 0 */
  int dim;                      /* embedding(0).dim(); 0 = no embedding */
  int threshold;                /* embedding(0).threshold(): grow V once fea_cnt > threshold */
  float v_lambda_l2;            /* embedding(0).lambda_l2() */
  int v_has_lr_eta;  float v_lr_eta;    /* embedding(0).lr_eta(), falls back to lr_eta when absent */
  int v_has_lr_beta; float v_lr_beta;
  float init_scale;             /* V starts uniform in [-init_scale, init_scale] */
  float grad_clipping, grad_normalization;
  float dropout;                /* embedding-gradient dropout; > 0 needs wh_difacto_set_dropout_stream */
  uint64_t key_begin, key_end;  /* the shard's key range (dense) */
  uint64_t max_key;             /* ps::FLAGS_max_key */
  uint64_t seed;                /* V init without an init stream: hash of (seed, key, j) */
} whDifactoConfig;

WH_API int wh_difacto_create(whContext* ctx, const whDifactoConfig* conf, whDifacto** out);
WH_API int wh_difacto_destroy(whDifacto* d);
/* The reference initialises a new embedding with glibc rand() in server arrival order
 * (learn/difacto/async_sgd.h:254). To reproduce a reference run, hand over the next n values of that
 * rand() sequence (host ints, after the same srand): the i-th embedding float created from now on
 * uses the i-th value. Without a stream V is initialised from a hash of (seed, key, j). */
WH_API int wh_difacto_set_init_stream(whDifacto* d, const int* rand_values, size_t n);
/* Gradient dropout (learn/difacto/loss.h:150-154) zeroes an embedding-gradient float when `(float)rand() / RAND_MAX >
 * 1 - dropout`, drawing from the WORKER process's glibc rand() -- one value per float of the keys that were pulled with
 * an embedding, ascending key order, every training minibatch. Hand over the next n values of that sequence; the i-th
 * float from now on uses the i-th value. shared_with_init != 0 (rand_values ignored): draw from the init stream and
 * advance its cursor instead -- worker and server living in ONE process share one rand() sequence (that is how the
 * checker runs the reference); in a ps-lite deployment they are separate processes with a sequence each. */
WH_API int wh_difacto_set_dropout_stream(whDifacto* d, const int* rand_values, size_t n, int shared_with_init);
/* AsyncWorker::ProcessMinibatch, learn/difacto/async_sgd.h:363-425: Localize with counts ->
 * (train, data_pass 0, dim > 0: count push, cmd = kPushFeaCnt) -> ZVPull -> Loss::Evaluate ->
 * Loss::CalcGrad -> ZVPush. prog[8] = {objv, auc, objv_w, copc(0), count, new_ex, new_w, new_V}
 * (learn/difacto/progress.h:28-36). Synchronous. */
WH_API int wh_difacto_process_minibatch(whDifacto* d, const whRowBlock* blk, int mem, int train, int data_pass,
                                        double* prog);
/* The same minibatch in phases, for key-range shards on several GPUs (rank g = AsyncWorker g + the AsyncServer shard of
 * [g*max_key/S, (g+1)*max_key/S)): ProcessMinibatch cut at its three exchanges -- the count push (learn/difacto/
 * async_sgd.h:374-381), ZVPull (:424), ZVPush (:408-411). All arrays are DEVICE pointers on the context's device.
 *   begin   Localize with counts (:366-371): the ascending unique keys and their counts (valid until the next begin).
 *   finish  Loss::Evaluate / CalcGrad (:385-407) with the pulled model in fixed-width rows -- w[U], has[U] (1: the key came
 *           with its embedding, model_siz = dim + 1; 0: w only), V[U x dim] -- writes gw[U], G[U x dim] for the push.
 *           prog[8] as wh_difacto_process_minibatch, new_w / new_V left 0 (the shards own them: shard_progress).
 *   shard_count_push / shard_pull / shard_push   what the server does with ONE message (:216-244): n keys of this shard's
 *           range, ascending. Messages are applied in the order of the calls. */
WH_API int wh_difacto_begin(whDifacto* d, const whRowBlock* blk, int mem, const uint64_t** d_uniq, const uint32_t** d_cnt,
                            size_t* num_uniq);
WH_API int wh_difacto_finish(whDifacto* d, const float* d_w, const unsigned char* d_has, const float* d_V, int train,
                             float* d_gw, float* d_G, double* prog);
WH_API int wh_difacto_shard_count_push(whDifacto* d, const uint64_t* d_keys, const uint32_t* d_cnt, size_t n);
WH_API int wh_difacto_shard_pull(whDifacto* d, const uint64_t* d_keys, size_t n, float* d_w, unsigned char* d_has, float* d_V);
WH_API int wh_difacto_shard_push(whDifacto* d, const uint64_t* d_keys, size_t n, const float* d_gw, const float* d_G,
                                 const unsigned char* d_has);
WH_API int wh_difacto_shard_progress(whDifacto* d, double* new_w, double* new_V);
/* ps::KVWorker<float> messages against this shard on HOST arrays -- what learn/difacto's worker sends (the C++ facade
 * host/ps.h forwards here). keys ascend. Synchronous.
 *   zpush_count  ZPush(feaid, feacnt, cmd = kPushFeaCnt), learn/difacto/async_sgd.h:374-379: counts are floats, added as
 *                (unsigned) counts[i] (:218).
 *   zvpull       ZVPull(feaid, val, val_siz), :424: sizes[i] = 1 ({w0}) or dim + 1 ({w0, V}), AdaGradHandle::Pull
 *                :234-244; vals has room for n * (dim + 1) floats, *vals_len = floats written.
 *   zvpush       ZVPush(feaid, val, val_siz), :408-411: the gradients in the layout zvpull returned. */
WH_API int wh_difacto_zpush_count(whDifacto* d, const uint64_t* keys, const float* counts, size_t n);
WH_API int wh_difacto_zvpull(whDifacto* d, const uint64_t* keys, size_t n, float* vals, int* sizes, size_t* vals_len);
WH_API int wh_difacto_zvpush(whDifacto* d, const uint64_t* keys, size_t n, const float* vals, const int* sizes, size_t vals_len);
/* Entry fields of the given keys (host arrays): fea_cnt[n], size[n] (1 or dim+1), w[n*(dim+1)] =
 * {w0, V}, cg[n*(dim+2)] = {sqc_grad_0, z_0, sqc_grad_V}; rows are zero beyond the entry's size. */
WH_API int wh_difacto_get_state(whDifacto* d, const uint64_t* keys, size_t n, uint32_t* fea_cnt, int* size,
                                float* w, float* cg);
/* difacto::Loss<float>(data, model, model_siz, conf); Evaluate; CalcGrad (learn/difacto/loss.h:24-158).
 * model is the ZVPull payload: model_siz[k] in {0, 1, dim+1} floats per key, concatenated; grad (NULL =
 * evaluate only) has the same layout. Host arrays. prog[8] as above. */
WH_API int wh_difacto_loss(whContext* ctx, const whLocalBlock* D, const float* model, size_t model_len,
                           const int* model_siz, size_t nkeys, int dim, float grad_clipping,
                           float grad_normalization, double* prog, float* grad);
/* The same with conf.dropout (learn/difacto/loss.h:150-154): rand_values = the caller's next rand() values, one per
 * float of model that belongs to an embedding (dim per key with model_siz = dim + 1), consumed in that order. */
WH_API int wh_difacto_loss_dropout(whContext* ctx, const whLocalBlock* D, const float* model, size_t model_len,
                                   const int* model_siz, size_t nkeys, int dim, float grad_clipping,
                                   float grad_normalization, float dropout, const int* rand_values, size_t num_rand,
                                   double* prog, float* grad);

#ifdef __cplusplus
}
#endif
#endif /* WORMHOLE_B200_H_ */
