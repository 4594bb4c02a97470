// Throughput of the atomic flavours the training step could use, per SM, on B200: build with
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_bench red_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ unsigned rnd(unsigned& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(unsigned long long* g64, unsigned* g32, unsigned nslots, int iters, int same) {
  extern __shared__ unsigned sm[];
  unsigned s = blockIdx.x * 7919u + threadIdx.x * 104729u + 1;
  for (int i = threadIdx.x; i < 16384; i += blockDim.x) sm[i] = 0;
  __syncthreads();
  for (int it = 0; it < iters; ++it) {
    unsigned r = rnd(s);
    unsigned slot = same ? (rnd(s) % same) : (r % nslots);     // same > 0: only `same` distinct hot addresses
    if (MODE == 0) atomicAdd(g64 + slot, (unsigned long long)r);
    if (MODE == 1) atomicAdd(g32 + slot, r);
    if (MODE == 2) atomicAdd(sm + (slot & 16383), r);
    if (MODE == 3) { unsigned old = atomicAdd(sm + (slot & 8191) * 2, r); if (old + r < old) atomicAdd(sm + (slot & 8191) * 2 + 1, 1u); }
    if (MODE == 4) sm[slot & 16383] = r;
    if (MODE == 5) { atomicAdd(g32 + slot * 2, r); atomicAdd(g32 + slot * 2 + 1, r); }
    if (MODE == 6) atomicOr(g32 + slot, 1u << (r & 31));
  }
  if (sm[threadIdx.x] == 0x12345) g32[0] = 1;
}

int main() {
  unsigned long long* g64; unsigned* g32;
  const unsigned big = 1u << 24;
  cudaMalloc(&g64, (size_t)big * 16); cudaMemset(g64, 0, (size_t)big * 16);
  g32 = (unsigned*)g64;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 2000, grid = 148, threads = 1024;
  const char* names[] = {"RED.64 global", "RED.32 global", "ATOMS.32 shared", "ATOMS 64-bit emulated (2x32)", "STS.32", "2x RED.32 global (adjacent)", "RED.OR.32 global"};
  auto run = [&](int mode, unsigned nslots, int same, const char* what) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      switch (mode) {
        case 0: k<0><<<grid, threads, 65536>>>(g64, g32, nslots, iters, same); break;
        case 1: k<1><<<grid, threads, 65536>>>(g64, g32, nslots, iters, same); break;
        case 2: k<2><<<grid, threads, 65536>>>(g64, g32, nslots, iters, same); break;
        case 3: k<3><<<grid, threads, 65536>>>(g64, g32, nslots, iters, same); break;
        case 4: k<4><<<grid, threads, 65536>>>(g64, g32, nslots, iters, same); break;
        case 5: k<5><<<grid, threads, 65536>>>(g64, g32, nslots, iters, same); break;
        case 6: k<6><<<grid, threads, 65536>>>(g64, g32, nslots, iters, same); break;
      }
      cudaEventRecord(e1); cudaEventSynchronize(e1);
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double ops = (double)grid * threads * iters;
    printf("%-30s %-28s %8.3f ms  %7.1f Gop/s  %.2f cyc/lane-op/SM (at 1.9 GHz)\n", names[mode], what, ms, ops / ms / 1e6,
           ms * 1e-3 * 1.9e9 / (threads * (double)iters));
  };
  cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  for (int mode : {0, 1, 5, 6}) {
    run(mode, 8192, 0, "8192 slots (L2 resident)");
    run(mode, 131072, 0, "128k slots");
    run(mode, big, 0, "16M slots (DRAM)");
    run(mode, 8192, 32, "32 hot addresses");
    run(mode, 8192, 1024, "1024 hot addresses");
  }
  for (int mode : {2, 3, 4}) { run(mode, 16384, 0, "spread"); run(mode, 16384, 32, "32 hot addresses"); }
  return 0;
}
