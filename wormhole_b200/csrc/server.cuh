// Piece (4): the HBM-resident model shard and its update handles. Internal API.
#pragma once
#include "common.cuh"
#include "spmv.cuh"

struct whServer {
  whContext* ctx = nullptr;
  int algo = WH_ALGO_FTRL;
  float l1 = 1.f, l2 = 0.f;
  float alpha = 0.1f, beta = 1.f;   // ISGDHandle defaults, learn/linear/async_sgd.h:49
  uint64_t key_begin = 0, key_end = 0;
  int t = 1;                        // SGDHandle::t, learn/linear/async_sgd.h:100
  wh::DevBuf table;                 // FTRL: float4 {w, z, sq_cum_grad, 0}; AdaGrad: float2 {w, sq_cum_grad}; SGD: float w
  wh::DevBuf counters;              // [0] new_w (long long), [1] out-of-range keys seen
  size_t entries() const { return (size_t)(key_end - key_begin); }
};

namespace wh {
// All take device pointers; d_n (may be NULL) overrides n with a count held on the device.
int server_pull_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, float* vals);
int server_push_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, const float* grads);
// Gradients still in the fixed-point accumulator of the transposed product: converts, clears the
// accumulator and applies the update in one pass (no fp32 gradient array in between).
int server_push_acc_launch(whServer* s, cudaStream_t stream, const uint64_t* keys, size_t n, const unsigned* d_n, long long* acc,
                           const FixedScale* fs, float mul, unsigned copies, size_t stride);
}  // namespace wh
