// The server's update handles as device functions: L1L2::Solve (learn/linear/penalty.h:36-41), the three handles of
// learn/linear/async_sgd.h (SGDHandle :83-102, AdaGradHandle :122-140, FTRLHandle :158-180) and ISGDHandle::Update
// (:35-41) on one entry of a shard (server.cuh: ShardView). Shared by the kernels of server.cu and direct.cu.
// All fp32 steps are single IEEE operations (-fmad=false; sqrtf and '/' are correctly rounded), so given the same
// gradient the new state is bit-identical to the reference's.
#pragma once
#include "server.cuh"

namespace wh {

// learn/linear/penalty.h:36-41
__device__ __forceinline__ float l1l2_solve(float z, float eta, float l1, float l2) {
  if (z <= l1 && z >= -l1) return 0.f;
  return (z > 0.f ? z - l1 : z + l1) / (eta + l2);
}

struct HandleParams {
  int algo;
  float l1, l2, alpha, beta;
  float eta_sgd;            // SGDHandle::eta for this push message
  uint64_t key_begin, key_end;
  // un-hashed 64-bit keys (ps::FLAGS_max_key at its default): open addressing instead of direct addressing
  unsigned long long* hkeys;   // NULL = dense shard; else one key per slot, kEmptyKey = free
  uint64_t hmask;              // slots - 1 (a power of two)
};
constexpr unsigned long long kEmptyKey = ~0ull;

__device__ __forceinline__ uint64_t hash_mix(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33; return x;
}
// Where does key k live? insert = true claims a free slot for a key not seen before (entries are zero by
// default, as ps-lite's store default-constructs on first touch); insert = false leaves absent keys absent.
// Returns false if the key is outside the shard / the table is full (counted as a bad key by the caller).
__device__ __forceinline__ bool find_slot(const HandleParams& h, uint64_t k, bool insert, uint64_t* slot, bool* present) {
  *present = true;
  if (h.hkeys == nullptr) {
    if (k < h.key_begin || k >= h.key_end) return false;
    *slot = k - h.key_begin;
    return true;
  }
  if (k == kEmptyKey) return false;
  uint64_t i = hash_mix(k) & h.hmask;
  for (uint64_t probes = 0; probes <= h.hmask; ++probes, i = (i + 1) & h.hmask) {
    unsigned long long cur = h.hkeys[i];
    if (cur == k) { *slot = i; return true; }
    if (cur == kEmptyKey) {
      if (!insert) { *present = false; *slot = i; return true; }
      cur = atomicCAS(h.hkeys + i, kEmptyKey, (unsigned long long)k);
      if (cur == kEmptyKey || cur == k) { *slot = i; return true; }
    }
  }
  return false;
}

// The handle's arithmetic for one key of one push message, on values in registers: takes {w, z, sq_cum_grad} and the
// gradient, leaves the new state in place and returns the new weight.
__device__ __forceinline__ float handle_update(const HandleParams& h, float old_w, float* z, float* n, float g) {
  if (h.algo == WH_ALGO_FTRL) {            // async_sgd.h:160-175
    const float sqrt_n = *n;
    const float n2 = sqrtf(sqrt_n * sqrt_n + g * g);
    const float sigma = (n2 - sqrt_n) / h.alpha;
    const float zz = *z + (g - sigma * old_w);
    *z = zz; *n = n2;
    return l1l2_solve(-zz, (h.beta + n2) / h.alpha, h.l1, h.l2);
  }
  if (h.algo == WH_ALGO_ADAGRAD) {         // async_sgd.h:123-135
    const float n0 = *n;
    const float n2 = sqrtf(n0 * n0 + g * g);
    const float eta = (n2 + h.beta) / h.alpha;
    *n = n2;
    return l1l2_solve(eta * old_w - g, eta, h.l1, h.l2);
  }
  return l1l2_solve(h.eta_sgd * old_w - g, h.eta_sgd, h.l1, h.l2);   // async_sgd.h:91-95
}
// ISGDHandle::Update (learn/linear/async_sgd.h:35-41): the change of |w|_0
__device__ __forceinline__ int nnz_delta(float old_w, float new_w) {
  return (old_w == 0.f && new_w != 0.f) ? 1 : ((old_w != 0.f && new_w == 0.f) ? -1 : 0);
}

// One key of one push message on the shard. Returns the change of |w|_0.
__device__ __forceinline__ int handle_push(const HandleParams& h, const ShardView& t, uint64_t slot, float g, float* new_w_out = nullptr) {
  float4* e = reinterpret_cast<float4*>(t.e + slot);     // {w, z, sq_cum_grad, -}: the first half of the entry
  float4 v = *e;
  const float old_w = v.x;
  const float new_w = handle_update(h, old_w, &v.y, &v.z, g);
  v.x = new_w;
  *e = v;
  if (new_w_out) *new_w_out = new_w;
  return nnz_delta(old_w, new_w);
}

__device__ __forceinline__ void commit_counters(long long* counters, int delta, int bad) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    delta += __shfl_xor_sync(0xffffffffu, delta, o);
    bad += __shfl_xor_sync(0xffffffffu, bad, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (delta) atomicAdd(reinterpret_cast<unsigned long long*>(counters), (unsigned long long)(long long)delta);
    if (bad) atomicAdd(reinterpret_cast<unsigned long long*>(counters + 1), (unsigned long long)bad);
  }
}

HandleParams make_params(whServer* s, bool start_push);

}  // namespace wh
