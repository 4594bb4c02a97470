// Piece (4): server-side sparse updates, applied in place to a dense direct-addressed table.
//
// Replaces the ps-lite server loop plus the three handles of learn/linear/async_sgd.h:
//   SGDHandle :83-102, AdaGradHandle :122-140, FTRLHandle :158-180, ISGDHandle::Update :35-41,
// with L1L2::Solve from learn/linear/penalty.h:36-41. One thread per key; entries are padded to
// 16 B (FTRL) / 8 B (AdaGrad) / 4 B (SGD) so that a key costs one vector load and one store.
// All fp32 steps are single IEEE operations (-fmad=false; sqrtf and '/' are correctly rounded),
// so given the same gradient the new state is bit-identical to the reference's.
#include <cooperative_groups.h>

#include <algorithm>
#include <vector>

#include "handle.cuh"
#include "scan.cuh"
#include "server.cuh"

namespace wh {

template <bool FROM_ACC>
__global__ void __launch_bounds__(256)
server_push_kernel(HandleParams h, ShardView table, const uint64_t* __restrict__ keys, size_t n,
                   const unsigned* d_n, const float* __restrict__ grads, long long* acc,
                   const FixedScale* fs, float mul, long long* counters, unsigned copies, size_t stride) {
  if (d_n) n = *d_n;
  int delta = 0, bad = 0;
  // grid-stride: the launch is sized for the number of SMs, not for the upper bound of n (which is only known
  // on the device and is ~15x smaller than the bound at Criteo shape)
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint64_t k = keys[i];
    uint64_t slot;
    bool present;
    const bool ok = find_slot(h, k, true, &slot, &present);
    // start the entry's trip from DRAM now, so that it overlaps the accumulator loads instead of following them
    if (ok) asm volatile("prefetch.global.L2 [%0];" ::"l"(table.e + slot));
    float g;
    if (FROM_ACC) {
      const long long a = sum_and_clear_replicas(acc, copies, stride, i);
      g = (float)((double)a * fs->inv) * mul;
    } else {
      g = grads[i];
    }
    if (ok) delta += handle_push(h, table, slot, g);
    else bad = 1;
  }
  commit_counters(counters, delta, bad);
}

// learn/linear/async_sgd.h:97-99,137-139,177-179: send[0] = w
__global__ void __launch_bounds__(256)
server_pull_kernel(HandleParams h, const ShardEntry* __restrict__ tab,
                   const uint64_t* __restrict__ keys, size_t n, const unsigned* d_n,
                   float* __restrict__ vals, long long* counters) {
  if (d_n) n = *d_n;
  int bad = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    uint64_t k = keys[i], s;
    bool present;
    if (find_slot(h, k, false, &s, &present)) vals[i] = present ? __ldg(&tab[s].w) : 0.f;
    else { vals[i] = 0.f; bad = 1; }
  }
  commit_counters(counters, 0, bad);
}

__global__ void server_state_kernel(HandleParams h, ShardView table,
                                    const uint64_t* __restrict__ keys, size_t n, float* w, float* z,
                                    float* cg, const float* sw, const float* sz, const float* scg,
                                    int set, long long* counters) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t k = keys[i], s;
  bool present;
  const int algo = h.algo;
  if (!find_slot(h, k, set != 0, &s, &present)) { atomicAdd(reinterpret_cast<unsigned long long*>(counters + 1), 1ull); return; }
  if (!present) { if (w) w[i] = 0.f; if (z) z[i] = 0.f; if (cg) cg[i] = 0.f; return; }
  float vw = table.e[s].w, vz = 0.f, vc = 0.f;
  if (algo == WH_ALGO_FTRL) { vz = table.e[s].z; vc = table.e[s].n; }
  else if (algo == WH_ALGO_ADAGRAD) vc = table.e[s].n;
  if (!set) { if (w) w[i] = vw; if (z) z[i] = vz; if (cg) cg[i] = vc; return; }
  float nw = sw ? sw[i] : vw, nz = sz ? sz[i] : vz, nc = scg ? scg[i] : vc;
  table.e[s].w = nw;
  if (algo == WH_ALGO_FTRL) { table.e[s].z = nz; table.e[s].n = nc; }
  else if (algo == WH_ALGO_ADAGRAD) table.e[s].n = nc;
  int delta = (vw == 0.f && nw != 0.f) ? 1 : ((vw != 0.f && nw == 0.f) ? -1 : 0);  // TLoad: Update(w, 0), :59-62
  if (delta) atomicAdd(reinterpret_cast<unsigned long long*>(counters), (unsigned long long)(long long)delta);
}

HandleParams make_params(whServer* s, bool start_push) {
  HandleParams h;
  h.algo = s->algo; h.l1 = s->l1; h.l2 = s->l2; h.alpha = s->alpha; h.beta = s->beta;
  h.key_begin = s->key_begin; h.key_end = s->key_end;
  h.hkeys = s->hkeys.as<unsigned long long>(); h.hmask = s->hslots ? s->hslots - 1 : 0;
  h.eta_sgd = 0.f;
  if (start_push) s->version++;
  if (start_push && s->algo == WH_ALGO_SGD) {  // SGDHandle::Start, async_sgd.h:85-90
    h.eta_sgd = (s->beta + sqrtf((float)s->t)) / s->alpha;
    s->t += 1;
  }
  return h;
}

int server_pull_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, float* vals) {
  if (n == 0) return WH_OK;
  WH_LAUNCH(server_pull_kernel, (unsigned)std::min<size_t>(div_up(n, 256), (size_t)s->ctx->num_sms * 8), 256, 0, stream, make_params(s, false), s->view().e,
            keys, n, d_n, vals, s->counters.as<long long>());
  return WH_OK;
}
int server_push_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, const float* grads) {
  HandleParams h = make_params(s, true);  // Start() runs even for an empty message
  if (n == 0) return WH_OK;
  WH_LAUNCH(server_push_kernel<false>, (unsigned)std::min<size_t>(div_up(n, 256), (size_t)s->ctx->num_sms * 8), 256, 0, stream, h, s->view(), keys, n,
            d_n, grads, (long long*)nullptr, (const FixedScale*)nullptr, 1.f, s->counters.as<long long>(), 1u, (size_t)0);
  return WH_OK;
}
int server_push_acc_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, long long* acc,
                           const FixedScale* fs, float mul, unsigned copies, size_t stride) {
  HandleParams h = make_params(s, true);
  if (n == 0) return WH_OK;
  WH_LAUNCH(server_push_kernel<true>, (unsigned)std::min<size_t>(div_up(n, 256), (size_t)s->ctx->num_sms * 8), 256, 0, stream, h, s->view(), keys, n,
            d_n, (const float*)nullptr, acc, fs, mul, s->counters.as<long long>(), copies, stride);
  return WH_OK;
}


// ---------------------------------------------------------------- shards on several GPUs --
// Rank g is worker g and server g (key range [bounds[g], bounds[g+1])). Instead of exchanging
// (key, value) messages, every rank maps every other rank's table (CUDA IPC; loads, stores and atomics
// go over NVLink) and runs the handle itself:
//   pull        one gather per key straight from the owner's table;
//   push        in `world` turns; in turn k rank r applies its gradients to shard (r + k) % world, so
//               each shard has exactly one writer at a time and every (worker, minibatch) gradient is
//               its own handle update, as in the reference's asynchronous server. Shard s receives its
//               sources in the order s, s-1, ..., s+1 (mod world);
//   barrier     between pull and the first turn and after every turn: a one-block kernel that announces
//               an epoch in every peer's flag array and waits for all peers' announcements. It is
//               stream-ordered -- no host synchronisation anywhere in the exchange.

__device__ __forceinline__ int shard_of(const ShardMap& m, uint64_t k) {
  int g = 0;
#pragma unroll
  for (int i = 1; i < kMaxShards; ++i) g += (i < m.world && k >= m.bounds[i]) ? 1 : 0;
  return g;
}

__global__ void __launch_bounds__(256)
server_pull_sharded_kernel(int algo, ShardMap m, uint64_t max_key, const uint64_t* __restrict__ keys, size_t n,
                           const unsigned* d_n, float* __restrict__ vals, long long* counters) {
  if (d_n) n = *d_n;
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int bad = 0;
  if (i < n) {
    const uint64_t k = keys[i];
    if (k < max_key) {
      const int g = shard_of(m, k);
      const uint64_t s = k - m.bounds[g];
      // ld.cg: the owner's L2 is the point of coherence for its table; a volatile load compiles to LDG.STRONG.SYS
      vals[i] = __ldcg(&static_cast<const ShardEntry*>(m.table[g])[s].w);
    } else { vals[i] = 0.f; bad = 1; }
  }
  commit_counters(counters, 0, bad);
}

__global__ void __launch_bounds__(256)
server_push_turn_kernel(HandleParams h, ShardMap m, int target, const uint64_t* __restrict__ keys, size_t n,
                        const unsigned* d_n, long long* acc, const FixedScale* fs, float mul, unsigned copies,
                        size_t stride) {
  if (d_n) n = *d_n;
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int delta = 0;
  if (i < n) {
    const uint64_t k = keys[i];
    if (k >= m.bounds[target] && k < m.bounds[target + 1]) {
      // all replicas first, then the clears: interleaved, every load waits for the one before it (the
      // stores may alias), and a thread pays `copies` memory latencies in a row
      long long a = sum_and_clear_replicas(acc, copies, stride, i);
      const float g = (float)((double)a * fs->inv) * mul;
      delta = handle_push(h, shard_view(m.table[target]), k - m.bounds[target], g);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) delta += __shfl_xor_sync(0xffffffffu, delta, o);
  if ((threadIdx.x & 31) == 0 && delta)
    atomicAdd_system(reinterpret_cast<unsigned long long*>(m.counters[target]), (unsigned long long)(long long)delta);
}

// All-rank rendezvous. flags[g][r] lives on GPU g and is written by rank r only.
// the barrier proper, for the first `world` threads of one block (callers fence before and after)
__device__ __forceinline__ void shard_barrier_wait(const ShardMap& m, unsigned long long epoch, long long* counters, int t, int cls = 0) {
  if (t < m.world) {
    volatile unsigned long long* out = m.flags[t] + cls * kMaxShards + m.rank;
    *out = epoch;
    __threadfence_system();
    const volatile unsigned long long* in = m.flags[m.rank] + cls * kMaxShards + t;
    const long long t0 = clock64();
    while (*in < epoch) {
      if (clock64() - t0 > 20000000000ll) {   // ~10 s: a peer is gone; report instead of hanging the GPU
        atomicAdd(reinterpret_cast<unsigned long long*>(counters + 2), 1ull);
        break;
      }
    }
  }
}

__global__ void shard_barrier_kernel(ShardMap m, unsigned long long epoch, long long* counters, int cls) {
  __threadfence_system();   // everything this GPU wrote to its peers before the barrier is visible first
  shard_barrier_wait(m, epoch, counters, threadIdx.x, cls);
  __threadfence_system();
}


// A worker's outbox: {header: U at word 0, slice cuts at words 4..4+world} | keys[cap] | grads[cap].
// The owner of a key range applies, for one source rank, the slice of that rank's (key, gradient) list that
// falls into its range -- read directly from the source GPU's memory. One launch per source, in ascending
// source order: each is one push message for this shard's handle.
// the gradient of one outbox entry: fp32 in the low half, or the fixed-point sum converted as fixed_finish_kernel does
__device__ __forceinline__ OutboxScale outbox_scale(const unsigned char* outbox) {
  OutboxScale sc;
  sc.inv = __ldcg(reinterpret_cast<const double*>(outbox + kOutboxScaleAt));
  sc.mul = __ldcg(reinterpret_cast<const float*>(outbox + kOutboxScaleAt + 8));
  sc.fixed = __ldcg(reinterpret_cast<const unsigned*>(outbox + kOutboxScaleAt + 12));
  return sc;
}
__device__ __forceinline__ float outbox_grad(long long g, const OutboxScale& sc) {
  return sc.fixed ? (float)((double)g * sc.inv) * sc.mul : __int_as_float((int)(unsigned)(unsigned long long)g);
}

__global__ void __launch_bounds__(256)
server_apply_from_kernel(HandleParams h, ShardView table, const unsigned char* outbox, size_t cap, int me, long long* counters) {
  // one thread per block fetches the two cut positions: they sit in the SOURCE GPU's memory, and tens of
  // thousands of threads loading the same remote word serialise on NVLink (measured: 8x slower push)
  __shared__ unsigned s_cut[2];
  __shared__ OutboxScale s_sc;
  if (threadIdx.x < 2) s_cut[threadIdx.x] = reinterpret_cast<const volatile unsigned*>(outbox)[4 + me + threadIdx.x];
  if (threadIdx.x == 2) s_sc = outbox_scale(outbox);
  __syncthreads();
  const unsigned lo = s_cut[0], hi = s_cut[1];
  const uint64_t* keys = reinterpret_cast<const uint64_t*>(outbox + outbox_keys_at());
  const long long* grads = reinterpret_cast<const long long*>(outbox + outbox_grads_at(cap));
  int delta = 0, bad = 0;
  for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (size_t)gridDim.x * blockDim.x) {
    const uint64_t k = keys[i];
    if (k >= h.key_begin && k < h.key_end) delta += handle_push(h, table, k - h.key_begin, outbox_grad(__ldcg(grads + i), s_sc));
    else bad = 1;
  }
  commit_counters(counters, delta, bad);
}

// The same `world` messages in ONE cooperative launch. Eight launches of ~30 k keys each were eight times a
// chain of {remote key read over NVLink -> local entry read -> update}, 288 us per push at 8 GPUs. Here every
// thread first issues its reads of ALL sources' slices (they are independent and overlap), then the sources are
// applied one after the other with a grid-wide barrier in between: the order per key -- ascending source rank,
// one message each, SGD's t advancing per message -- is exactly the one of the separate launches.
struct ApplyAllArgs {
  const unsigned char* outbox[kMaxShards];
  float eta_sgd[kMaxShards];
  int world;
};
__global__ void __launch_bounds__(256)
server_apply_all_kernel(HandleParams h, ShardView table, ApplyAllArgs a, size_t cap, int me, long long* counters) {
  // (the barriers before and after stay separate one-block launches: folded into this kernel they kept a block on
  //  every SM spinning while the peers caught up -- 4.36 -> 3.92 G examples/s at 8 GPUs)
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  __shared__ unsigned s_cut[kMaxShards][2];
  __shared__ OutboxScale s_sc[kMaxShards];
  if (threadIdx.x < 2 * a.world) {   // one thread per cut: the headers sit in the source GPUs' memory
    const int src = threadIdx.x >> 1;
    s_cut[src][threadIdx.x & 1] = reinterpret_cast<const volatile unsigned*>(a.outbox[src])[4 + me + (threadIdx.x & 1)];
  } else if (threadIdx.x >= 32 && threadIdx.x < 32 + a.world) {
    const int src = threadIdx.x - 32;
    s_sc[src] = outbox_scale(a.outbox[src]);
  }
  __syncthreads();
  const size_t gt = (size_t)blockIdx.x * blockDim.x + threadIdx.x, gsz = (size_t)gridDim.x * blockDim.x;
  uint64_t k0[kMaxShards];
  long long g0[kMaxShards];
#pragma unroll
  for (int src = 0; src < kMaxShards; ++src) {
    if (src < a.world) {
      const size_t i = s_cut[src][0] + gt;
      if (i < s_cut[src][1]) {
        k0[src] = __ldcg(reinterpret_cast<const uint64_t*>(a.outbox[src] + outbox_keys_at()) + i);
        g0[src] = __ldcg(reinterpret_cast<const long long*>(a.outbox[src] + outbox_grads_at(cap)) + i);
      }
    }
  }
  int delta = 0, bad = 0;
#pragma unroll
  for (int src = 0; src < kMaxShards; ++src) {
    if (src < a.world) {
      h.eta_sgd = a.eta_sgd[src];
      const uint64_t* keys = reinterpret_cast<const uint64_t*>(a.outbox[src] + outbox_keys_at());
      const long long* grads = reinterpret_cast<const long long*>(a.outbox[src] + outbox_grads_at(cap));
      const size_t lo = s_cut[src][0], hi = s_cut[src][1];
      size_t i = lo + gt;
      if (i < hi) {
        if (k0[src] >= h.key_begin && k0[src] < h.key_end) delta += handle_push(h, table, k0[src] - h.key_begin, outbox_grad(g0[src], s_sc[src]));
        else bad = 1;
        for (i += gsz; i < hi; i += gsz) {   // slices longer than the grid
          const uint64_t k = __ldcg(keys + i);
          if (k >= h.key_begin && k < h.key_end) delta += handle_push(h, table, k - h.key_begin, outbox_grad(__ldcg(grads + i), s_sc[src]));
          else bad = 1;
        }
      }
      if (src + 1 < a.world) grid.sync();
    }
  }
  commit_counters(counters, delta, bad);
}

// Blocks per SM of the two exchange kernels (256 threads each). More threads = fewer trips per thread, and every trip is
// a chain of a remote read and a local access; but the apply kernel is cooperative -- all its blocks must be resident
// at once, beside whatever the other minibatches in flight are running -- and stalls when it asks for too much
// (measured on 2 GPUs, ms per step: apply 1 / 2 / 4 blocks per SM = 0.150 / 0.143 / 0.302).
// Start of a lane's next minibatch: the gradient words of the previous one (U is still in the header) back to zero --
// the step's kernel adds its fixed-point sums into them
__global__ void outbox_clear_kernel(unsigned char* outbox, size_t cap) {
  const unsigned U = reinterpret_cast<const unsigned*>(outbox)[0];
  long long* g = reinterpret_cast<long long*>(outbox + outbox_grads_at(cap));
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < U && i < cap; i += (size_t)gridDim.x * blockDim.x) g[i] = 0;
}
int outbox_clear_launch(whServer* s, cudaStream_t stream, unsigned char* outbox, size_t cap) {
  WH_LAUNCH(outbox_clear_kernel, (unsigned)s->ctx->num_sms * 2, 256, 0, stream, outbox, cap);
  return WH_OK;
}

static unsigned exchange_blocks_per_sm(bool cooperative) {
  const char* e = getenv(cooperative ? "WH_APPLY_BLOCKS" : "WH_SERVE_BLOCKS");
  const int b = e ? atoi(e) : (cooperative ? 2 : 4);
  return (unsigned)(b < 1 ? 1 : (b > 8 ? 8 : b));
}

// cuts[g] = first position in the ascending key list whose key is >= bounds[g]
__global__ void outbox_cuts_kernel(unsigned char* outbox, ShardMap m) {
  unsigned* hdr = reinterpret_cast<unsigned*>(outbox);
  const uint64_t* keys = reinterpret_cast<const uint64_t*>(outbox + 256);
  const unsigned U = hdr[0];
  const int g = threadIdx.x;
  if (g > m.world) return;
  unsigned lo = 0, hi = U;
  const uint64_t b = m.bounds[g];
  while (lo < hi) { unsigned mid = (lo + hi) >> 1; if (keys[mid] < b) lo = mid + 1; else hi = mid; }
  hdr[4 + g] = (g == m.world) ? U : lo;
}

int server_apply_from_launch(whServer* s, cudaStream_t stream, const unsigned char* outbox, size_t cap) {
  HandleParams h = make_params(s, true);   // one push message: SGD's t advances even for an empty slice
  WH_LAUNCH(server_apply_from_kernel, (unsigned)s->ctx->num_sms, 256, 0, stream, h, s->view(), outbox, cap,
            s->map.rank, s->counters.as<long long>());
  return WH_OK;
}
int server_apply_all_launch(whServer* s, cudaStream_t stream, unsigned char* const* outboxes, size_t cap) {
  ApplyAllArgs a;
  a.world = s->map.world;
  HandleParams h = make_params(s, false);
  for (int src = 0; src < a.world; ++src) {
    a.outbox[src] = outboxes[src];
    // one push message per source that has a minibatch in this step: SGDHandle::Start each
    a.eta_sgd[src] = ((s->active_sources >> src) & 1) ? make_params(s, true).eta_sgd : 0.f;
  }
  int me = s->map.rank;
  ShardView table = s->view();
  long long* counters = s->counters.as<long long>();
  void* args[] = {&h, &table, &a, &cap, &me, &counters};
  const unsigned grid = (unsigned)s->ctx->num_sms * exchange_blocks_per_sm(true);
  WH_CUDA(cudaLaunchCooperativeKernel((const void*)server_apply_all_kernel, dim3(grid), dim3(256), args, 0, stream));
  wh::g_launches.fetch_add(1, std::memory_order_relaxed);
  return WH_OK;
}
int outbox_cuts_launch(whServer* s, cudaStream_t stream, unsigned char* outbox) {
  WH_LAUNCH(outbox_cuts_kernel, 1, 32, 0, stream, outbox, s->map);
  return WH_OK;
}

int server_pull_sharded_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, float* vals) {
  if (n == 0) return WH_OK;
  WH_LAUNCH(server_pull_sharded_kernel, (unsigned)div_up(n, 256), 256, 0, stream, s->algo, s->map,
            s->map.bounds[s->map.world], keys, n, d_n, vals, s->counters.as<long long>());
  return WH_OK;
}

int server_push_turn_launch(whServer* s, cudaStream_t stream, int turn, const uint64_t* keys, size_t n, const unsigned* d_n,
                            long long* acc, const FixedScale* fs, float mul, unsigned copies, size_t stride) {
  const int target = (s->map.rank + turn) % s->map.world;
  HandleParams h;
  h.algo = s->algo; h.l1 = s->l1; h.l2 = s->l2; h.alpha = s->alpha; h.beta = s->beta;
  h.key_begin = s->map.bounds[target]; h.key_end = s->map.bounds[target + 1];
  h.hkeys = nullptr; h.hmask = 0;
  h.eta_sgd = 0.f;
  if (s->algo == WH_ALGO_SGD) {
    // SGDHandle::t of the target shard when this message arrives: it has seen `world` messages per earlier
    // step and `turn` messages in this one (learn/linear/async_sgd.h:85-90)
    const unsigned long long t = 1ull + s->steps * (unsigned long long)s->map.world + (unsigned long long)turn;
    h.eta_sgd = (s->beta + sqrtf((float)t)) / s->alpha;
  }
  if (n == 0) return WH_OK;
  WH_LAUNCH(server_push_turn_kernel, (unsigned)div_up(n, 256), 256, 0, stream, h, s->map, target, keys, n, d_n, acc, fs,
            mul, copies, stride);
  return WH_OK;
}

int server_barrier_launch(whServer* s, cudaStream_t stream, int cls) {
  if (s->map.world <= 1) return WH_OK;
  // flags only ever grow, so the barriers of one class must complete in the order they were issued, whichever
  // streams they sit in: each waits for its predecessor of the same class
  if (!s->barrier_ev[cls]) WH_CUDA(cudaEventCreateWithFlags(&s->barrier_ev[cls], cudaEventDisableTiming));
  else WH_CUDA(cudaStreamWaitEvent(stream, s->barrier_ev[cls], 0));
  s->epoch[cls]++;
  WH_LAUNCH(shard_barrier_kernel, 1, 32, 0, stream, s->map, s->epoch[cls], s->counters.as<long long>(), cls);
  WH_CUDA(cudaEventRecord(s->barrier_ev[cls], stream));
  return WH_OK;
}

// ---- owner-served pull ---------------------------------------------------------------------------------------
// 224 k scattered 4-byte reads over NVLink per pull (one per remote key) took 42 us at 8 GPUs: the fabric is bound by
// the number of requests, not by bytes. Here the OWNER reads each source's slice of keys (contiguous: coalesced
// 128-byte requests), gathers the weights from its own HBM and writes them into the source's inbox, again coalesced.
struct ServeArgs {
  unsigned char* outbox[kMaxShards];
  int world;
};
__global__ void __launch_bounds__(256)
server_serve_all_kernel(const ShardEntry* tab, uint64_t key_begin, uint64_t key_end, ServeArgs a, size_t cap, int me,
                        long long* counters) {
  __shared__ unsigned s_cut[kMaxShards][2];
  if (threadIdx.x < 2 * a.world) {   // one thread per cut: the headers sit in the source GPUs' memory
    const int src = threadIdx.x >> 1;
    s_cut[src][threadIdx.x & 1] = reinterpret_cast<const volatile unsigned*>(a.outbox[src])[4 + me + (threadIdx.x & 1)];
  }
  __syncthreads();
  const size_t gt = (size_t)blockIdx.x * blockDim.x + threadIdx.x, gsz = (size_t)gridDim.x * blockDim.x;
  int bad = 0;
  for (size_t i0 = gt; ; i0 += gsz) {
    // one position of every source's slice per trip: the remote key reads of a trip are independent
    uint64_t k[kMaxShards];
    bool live[kMaxShards], any = false;
#pragma unroll
    for (int src = 0; src < kMaxShards; ++src) {
      live[src] = src < a.world && s_cut[src][0] + i0 < s_cut[src][1];
      if (live[src]) { k[src] = __ldcg(reinterpret_cast<const uint64_t*>(a.outbox[src] + 256) + s_cut[src][0] + i0); any = true; }
    }
    if (!__syncthreads_or(any)) break;
#pragma unroll
    for (int src = 0; src < kMaxShards; ++src) {
      if (!live[src]) continue;
      float w = 0.f;
      if (k[src] >= key_begin && k[src] < key_end) w = __ldcg(&tab[k[src] - key_begin].w);
      else bad = 1;
      reinterpret_cast<float*>(a.outbox[src] + outbox_inbox_at(cap))[s_cut[src][0] + i0] = w;
    }
  }
  commit_counters(counters, 0, bad);
}

int server_serve_all_launch(whServer* s, cudaStream_t stream, unsigned char* const* outboxes, size_t cap) {
  ServeArgs a;
  a.world = s->map.world;
  for (int src = 0; src < a.world; ++src) a.outbox[src] = outboxes[src];
  WH_LAUNCH(server_serve_all_kernel, (unsigned)s->ctx->num_sms * exchange_blocks_per_sm(false), 256, 0, stream, s->view().e, s->key_begin, s->key_end, a, cap,
            s->map.rank, s->counters.as<long long>());
  return WH_OK;
}


// ---- model files (the train -> save -> load -> predict loop) ------------------------------------------
// Entries save only w, and entries with w == 0 are skipped (SGDEntry/AdaGradEntry/FTRLEntry::Save and ::Empty,
// learn/linear/async_sgd.h:59-66,73-75,111-113,150-152). The record framing of ps-lite's KVStore::Save is
// not in the reference tree; this library writes {uint64 key, float w} records in ascending key order.
struct NonZero {
  const ShardEntry* tab;
  __device__ float w(size_t i) const { return tab[i].w; }
  __device__ unsigned operator()(size_t i) const { return w(i) != 0.f ? 1u : 0u; }
};
struct DumpOut {
  NonZero nz; uint64_t key_begin; const unsigned long long* hkeys; uint64_t* keys; float* ws;
  __device__ void operator()(size_t i, unsigned v, unsigned prefix) const {
    if (v) { keys[prefix] = hkeys ? hkeys[i] : key_begin + i; ws[prefix] = nz.w(i); }
  }
};

}  // namespace wh

using namespace wh;

static int check_bad_keys(whServer* s, const char* what) {
  whContext* ctx = s->ctx;
  WH_CUDA(cudaMemcpyAsync(ctx->pinned.p, s->counters.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  long long bad = ctx->pinned.as<long long>()[1];
  if (bad) {
    WH_CUDA(cudaMemsetAsync(s->counters.as<long long>() + 1, 0, 8, ctx->stream));
    set_error("%s: %lld keys outside this shard's range [%llu, %llu)", what, bad,
              (unsigned long long)s->key_begin, (unsigned long long)s->key_end);
    return WH_ERR_INVALID;
  }
  return WH_OK;
}

extern "C" {

int wh_server_create(whContext* ctx, const whServerConfig* c, whServer** out) {
  WH_REQUIRE(ctx && c && out, "wh_server_create: NULL argument");
  WH_REQUIRE(c->algo == WH_ALGO_SGD || c->algo == WH_ALGO_ADAGRAD || c->algo == WH_ALGO_FTRL,
             "wh_server_create: unknown algo: %d", c->algo);
  WH_REQUIRE(c->lambda_l1 >= 0 && c->lambda_l2 >= 0, "wh_server_create: lambda must be >= 0 (learn/linear/penalty.h:15,19)");
  WH_REQUIRE(c->key_end > c->key_begin, "wh_server_create: empty key range");
  WH_REQUIRE(c->hash_slots != 0 || c->key_end - c->key_begin <= (1ull << 32),
             "wh_server_create: a dense shard of %llu keys is too large; give hash_slots for un-hashed 64-bit keys",
             (unsigned long long)(c->key_end - c->key_begin));
  WH_REQUIRE((c->hash_slots & (c->hash_slots - 1)) == 0 && c->hash_slots <= (1ull << 32), "wh_server_create: hash_slots must be a power of two <= 2^32");
  DeviceGuard g(ctx->device);
  whServer* s = new whServer();
  s->ctx = ctx; s->algo = c->algo; s->l1 = c->lambda_l1; s->l2 = c->lambda_l2;
  if (c->has_lr_eta) s->alpha = c->lr_eta;
  if (c->has_lr_beta) s->beta = c->lr_beta;
  s->key_begin = c->key_begin; s->key_end = c->key_end;
  s->hslots = c->hash_slots;
  int r = s->table.reserve(shard_bytes(s->entries()));
  if (r == WH_OK && s->hslots) r = s->hkeys.reserve(s->hslots * 8);
  if (r == WH_OK) r = s->counters.reserve(64);
  if (r == WH_OK) r = s->flags.reserve(kBarrierClasses * kMaxShards * 8);
  if (r != WH_OK) { delete s; return r; }
  WH_CUDA(cudaMemsetAsync(s->table.p, 0, s->table.cap, ctx->stream));  // entries default to 0 (:72,108-109,146-148)
  WH_CUDA(cudaMemsetAsync(s->counters.p, 0, 64, ctx->stream));
  WH_CUDA(cudaMemsetAsync(s->flags.p, 0, kBarrierClasses * kMaxShards * 8, ctx->stream));
  if (s->hslots) WH_CUDA(cudaMemsetAsync(s->hkeys.p, 0xff, s->hslots * 8, ctx->stream));
  s->map.world = 1; s->map.rank = 0;
  s->map.bounds[0] = c->key_begin; s->map.bounds[1] = c->key_end;
  s->map.table[0] = s->table.p; s->map.counters[0] = s->counters.as<long long>();
  s->map.flags[0] = s->flags.as<unsigned long long>();
  *out = s;
  return WH_OK;
}

int wh_server_destroy(whServer* s) {
  if (!s) return WH_OK;
  DeviceGuard g(s->ctx->device);
  cudaStreamSynchronize(s->ctx->stream);
  for (void* p : s->peer_base) if (p) cudaIpcCloseMemHandle(p);
  for (auto e : s->barrier_ev) if (e) cudaEventDestroy(e);
  s->table.release(); s->counters.release(); s->flags.release(); s->hkeys.release();
  delete s;
  return WH_OK;
}

int wh_server_pull(whServer* s, const uint64_t* keys, size_t n, int mem, float* vals) {
  WH_REQUIRE(s && (n == 0 || (keys && vals)), "wh_server_pull: NULL argument");
  whContext* ctx = s->ctx;
  DeviceGuard g(ctx->device);
  const uint64_t* dk; float* dv;
  WH_TRY(to_device(ctx, ctx->stage[0], keys, n, mem, &dk));
  WH_TRY(out_buffer(ctx, ctx->stage[1], vals, n, mem, &dv));
  WH_TRY(server_pull_launch(s, ctx->stream, dk, n, nullptr, dv));
  WH_TRY(out_commit(ctx, vals, dv, n, mem));
  if (mem == WH_HOST) return check_bad_keys(s, "wh_server_pull");
  return WH_OK;
}

int wh_server_push(whServer* s, const uint64_t* keys, size_t n, int mem, const float* grads) {
  WH_REQUIRE(s && (n == 0 || (keys && grads)), "wh_server_push: NULL argument");
  whContext* ctx = s->ctx;
  DeviceGuard g(ctx->device);
  const uint64_t* dk; const float* dg;
  WH_TRY(to_device(ctx, ctx->stage[0], keys, n, mem, &dk));
  WH_TRY(to_device(ctx, ctx->stage[1], grads, n, mem, &dg));
  WH_TRY(server_push_launch(s, ctx->stream, dk, n, nullptr, dg));
  if (mem == WH_HOST) return check_bad_keys(s, "wh_server_push");
  return WH_OK;
}

int wh_server_nnz_w(whServer* s, int64_t* out) {
  WH_REQUIRE(s && out, "wh_server_nnz_w: NULL argument");
  whContext* ctx = s->ctx;
  DeviceGuard g(ctx->device);
  WH_CUDA(cudaMemcpyAsync(ctx->pinned.p, s->counters.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  *out = ctx->pinned.as<long long>()[0];
  return WH_OK;
}

static int state_io(whServer* s, const uint64_t* keys, size_t n, int mem, float* w, float* z, float* cg,
                    const float* sw, const float* sz, const float* scg, int set) {
  whContext* ctx = s->ctx;
  DeviceGuard g(ctx->device);
  if (n == 0) return WH_OK;
  const uint64_t* dk; float *dw = nullptr, *dz = nullptr, *dc = nullptr; const float *dsw, *dsz, *dsc;
  WH_TRY(to_device(ctx, ctx->stage[0], keys, n, mem, &dk));
  if (w) WH_TRY(out_buffer(ctx, ctx->stage[1], w, n, mem, &dw));
  if (z) WH_TRY(out_buffer(ctx, ctx->stage[2], z, n, mem, &dz));
  if (cg) WH_TRY(out_buffer(ctx, ctx->stage[3], cg, n, mem, &dc));
  WH_TRY(to_device(ctx, ctx->stage[4], sw, n, mem, &dsw));
  WH_TRY(to_device(ctx, ctx->stage[5], sz, n, mem, &dsz));
  WH_TRY(to_device(ctx, ctx->stage[6], scg, n, mem, &dsc));
  if (set) s->version++;
  WH_LAUNCH(server_state_kernel, (unsigned)div_up(n, 256), 256, 0, ctx->stream, make_params(s, false), s->view(), dk, n, dw, dz, dc, dsw, dsz, dsc, set, s->counters.as<long long>());
  WH_TRY(out_commit(ctx, w, dw, n, mem));
  WH_TRY(out_commit(ctx, z, dz, n, mem));
  WH_TRY(out_commit(ctx, cg, dc, n, mem));
  if (mem == WH_HOST) return check_bad_keys(s, set ? "wh_server_set_state" : "wh_server_get_state");
  return WH_OK;
}

int wh_server_get_state(whServer* s, const uint64_t* keys, size_t n, int mem, float* w, float* z, float* cg) {
  WH_REQUIRE(s && (n == 0 || keys), "wh_server_get_state: NULL argument");
  return state_io(s, keys, n, mem, w, z, cg, nullptr, nullptr, nullptr, 0);
}
int wh_server_set_state(whServer* s, const uint64_t* keys, size_t n, int mem, const float* w, const float* z,
                        const float* cg) {
  WH_REQUIRE(s && (n == 0 || keys), "wh_server_set_state: NULL argument");
  return state_io(s, keys, n, mem, nullptr, nullptr, nullptr, w, z, cg, 1);
}
int wh_server_sgd_t(const whServer* s) {
  if (!s) return 0;
  if (s->map.world > 1) return (int)(1 + s->steps * (unsigned long long)s->map.world);   // lock-step shards
  return s->t;
}


int wh_server_save(whServer* s, const char* path) {
  WH_REQUIRE(s && path, "wh_server_save: NULL argument");
  whContext* ctx = s->ctx;
  DeviceGuard g(ctx->device);
  int64_t nnz = 0;
  WH_TRY(wh_server_nnz_w(s, &nnz));
  const size_t n = s->entries(), cap = (size_t)(nnz > 0 ? nnz : 0) + 1;
  WH_TRY(ctx->stage[0].reserve(cap * 8)); WH_TRY(ctx->stage[1].reserve(cap * 4));
  WH_TRY(ctx->scan_tmp.reserve(div_up(n, kScanTile) * sizeof(unsigned) + 256));
  unsigned* total = ctx->small.as<unsigned>() + 64;
  NonZero nz{s->view().e};
  WH_TRY(scan_and_apply<unsigned>(ctx, nz, DumpOut{nz, s->key_begin, s->hkeys.as<unsigned long long>(), ctx->stage[0].as<uint64_t>(), ctx->stage[1].as<float>()},
                                  n, ctx->scan_tmp.as<unsigned>(), total));
  std::vector<uint64_t> keys(cap);
  std::vector<float> ws(cap);
  unsigned cnt = 0;
  WH_CUDA(cudaMemcpyAsync(&cnt, total, 4, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaMemcpyAsync(keys.data(), ctx->stage[0].p, cap * 8, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaMemcpyAsync(ws.data(), ctx->stage[1].p, cap * 4, cudaMemcpyDeviceToHost, ctx->stream));
  WH_CUDA(cudaStreamSynchronize(ctx->stream));
  WH_REQUIRE((int64_t)cnt == nnz, "wh_server_save: %u non-zero weights found but |w|_0 counter says %lld", cnt, (long long)nnz);
  std::vector<unsigned> order(cnt);
  for (unsigned i = 0; i < cnt; ++i) order[i] = i;
  if (s->hslots) std::sort(order.begin(), order.end(), [&](unsigned a, unsigned b) { return keys[a] < keys[b]; });  // slot order -> key order
  FILE* f = fopen(path, "wb");
  WH_REQUIRE(f != nullptr, "wh_server_save: cannot open %s", path);
  bool ok = true;
  for (unsigned i = 0; i < cnt && ok; ++i) ok = fwrite(&keys[order[i]], 8, 1, f) == 1 && fwrite(&ws[order[i]], 4, 1, f) == 1;
  ok = (fclose(f) == 0) && ok;
  WH_REQUIRE(ok, "wh_server_save: writing %s failed (disk full?)", path);
  return WH_OK;
}

int wh_server_load(whServer* s, const char* path) {
  WH_REQUIRE(s && path, "wh_server_load: NULL argument");
  FILE* f = fopen(path, "rb");
  WH_REQUIRE(f != nullptr, "wh_server_load: cannot open %s", path);
  std::vector<uint64_t> keys; std::vector<float> ws;
  uint64_t k; float w;
  while (fread(&k, 8, 1, f) == 1 && fread(&w, 4, 1, f) == 1) { keys.push_back(k); ws.push_back(w); }
  fclose(f);
  // TLoad: only w is read back, and Update(w, 0) counts it (learn/linear/async_sgd.h:59-62); z and
  // sq_cum_grad of a loaded entry stay at their defaults, as in the reference
  return wh_server_set_state(s, keys.data(), keys.size(), WH_HOST, ws.data(), nullptr, nullptr);
}

int wh_server_ipc_export(whServer* s, void* handle) {
  WH_REQUIRE(s && handle, "wh_server_ipc_export: NULL argument");
  DeviceGuard g(s->ctx->device);
  WH_CUDA(cudaStreamSynchronize(s->ctx->stream));
  cudaIpcMemHandle_t* h = reinterpret_cast<cudaIpcMemHandle_t*>(handle);
  WH_CUDA(cudaIpcGetMemHandle(&h[0], s->table.p));
  WH_CUDA(cudaIpcGetMemHandle(&h[1], s->counters.p));
  WH_CUDA(cudaIpcGetMemHandle(&h[2], s->flags.p));
  return WH_OK;
}

int wh_server_ipc_attach(whServer* s, int rank, int world, uint64_t max_key, const void* handles) {
  WH_REQUIRE(s && handles, "wh_server_ipc_attach: NULL argument");
  WH_REQUIRE(world >= 1 && world <= kMaxShards && rank >= 0 && rank < world, "wh_server_ipc_attach: rank %d of %d", rank, world);
  WH_REQUIRE(s->map.world == 1, "wh_server_ipc_attach: already attached");
  WH_REQUIRE(s->hslots == 0, "wh_server_ipc_attach: key-range shards over several GPUs need a hashed key space (dense shards)");
  const uint64_t b = (uint64_t)(((unsigned __int128)rank * max_key) / world), e = (uint64_t)(((unsigned __int128)(rank + 1) * max_key) / world);
  WH_REQUIRE(s->key_begin == b && s->key_end == e, "wh_server_ipc_attach: this shard holds [%llu, %llu) but rank %d of %d owns [%llu, %llu)",
             (unsigned long long)s->key_begin, (unsigned long long)s->key_end, rank, world, (unsigned long long)b, (unsigned long long)e);
  DeviceGuard g(s->ctx->device);
  const cudaIpcMemHandle_t* h = reinterpret_cast<const cudaIpcMemHandle_t*>(handles);
  ShardMap m;
  m.world = world; m.rank = rank;
  for (int r = 0; r <= world; ++r) m.bounds[r] = (uint64_t)(((unsigned __int128)r * max_key) / world);
  for (int r = 0; r < world; ++r) {
    if (r == rank) {
      m.table[r] = s->table.p; m.counters[r] = s->counters.as<long long>(); m.flags[r] = s->flags.as<unsigned long long>();
      continue;
    }
    void* p[3] = {nullptr, nullptr, nullptr};
    for (int j = 0; j < 3; ++j) {
      WH_CUDA(cudaIpcOpenMemHandle(&p[j], h[3 * r + j], cudaIpcMemLazyEnablePeerAccess));
      s->peer_base[3 * r + j] = p[j];
    }
    m.table[r] = p[0]; m.counters[r] = static_cast<long long*>(p[1]); m.flags[r] = static_cast<unsigned long long*>(p[2]);
  }
  s->map = m;
  return WH_OK;
}

}  // extern "C"
