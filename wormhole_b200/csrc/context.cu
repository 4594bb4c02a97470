// Context life cycle, error strings, scratch buffers.
#include <stdarg.h>

#include <stdlib.h>

#include <string>

#include "common.cuh"

namespace wh {

static thread_local char g_err[1024] = "";
std::atomic<uint64_t> g_launches{0};
FILE* g_timeline = nullptr;
static cudaEvent_t g_tl_base = nullptr;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int DevBuf::reserve(size_t bytes) {
  if (bytes <= cap) return WH_OK;
  if (p) { WH_CUDA(cudaFree(p)); p = nullptr; cap = 0; }
  size_t want = align_up(bytes + bytes / 8, 256);
  cudaError_t e = cudaMalloc(&p, want);
  if (e != cudaSuccess) {
    p = nullptr;
    set_error("cudaMalloc(%zu bytes) failed: %s", want, cudaGetErrorString(e));
    return WH_ERR_NOMEM;
  }
  cap = want;
  return WH_OK;
}
void DevBuf::release() { if (p) cudaFree(p); p = nullptr; cap = 0; }

int PinnedBuf::reserve(size_t bytes) {
  if (bytes <= cap) return WH_OK;
  if (p) { WH_CUDA(cudaFreeHost(p)); p = nullptr; cap = 0; }
  size_t want = align_up(bytes, 4096);
  cudaError_t e = cudaMallocHost(&p, want);
  if (e != cudaSuccess) {
    p = nullptr;
    set_error("cudaMallocHost(%zu bytes) failed: %s", want, cudaGetErrorString(e));
    return WH_ERR_NOMEM;
  }
  cap = want;
  return WH_OK;
}
void PinnedBuf::release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }

int prof_begin(whContext* ctx) {
  if (!ctx->prof.on) return WH_OK;
  return prof_mark(ctx, nullptr);
}
int prof_mark(whContext* ctx, const char* name) {
  Profiler& p = ctx->prof;
  if (p.used == p.ev.size()) {
    cudaEvent_t e;
    WH_CUDA(cudaEventCreate(&e));
    p.ev.push_back(e);
    p.name.push_back(nullptr);
  }
  if (g_timeline && !g_tl_base) { WH_CUDA(cudaEventCreate(&g_tl_base)); WH_CUDA(cudaEventRecord(g_tl_base, ctx->stream)); }
  WH_CUDA(cudaEventRecord(p.ev[p.used], ctx->stream));
  p.name[p.used] = name;  // name of the segment that ENDS at this event (NULL = start marker)
  p.used++;
  return WH_OK;
}
int prof_collect(whContext* ctx) {
  Profiler& p = ctx->prof;
  if (g_timeline)
    for (size_t i = 0; i < p.used; ++i) {
      float t = 0;
      if (cudaEventElapsedTime(&t, g_tl_base, p.ev[i]) == cudaSuccess)
        fprintf(g_timeline, "%p %-12s %10.1f\n", (void*)ctx, p.name[i] ? p.name[i] : "begin", (double)t * 1e3);
    }
  for (size_t i = 1; i < p.used; ++i) {
    if (!p.name[i]) continue;
    float ms = 0;
    WH_CUDA(cudaEventElapsedTime(&ms, p.ev[i - 1], p.ev[i]));
    Profiler::Acc* a = nullptr;
    for (auto& x : p.acc) if (x.name == p.name[i]) { a = &x; break; }
    if (!a) { p.acc.push_back(Profiler::Acc()); a = &p.acc.back(); a->name = p.name[i]; }
    a->ms += ms; a->count++;
  }
  p.used = 0;
  return WH_OK;
}

}  // namespace wh

extern "C" {

int wh_profile_enable(whContext* c, int on) {
  WH_REQUIRE(c != nullptr, "wh_profile_enable: ctx is NULL");
  wh::DeviceGuard g(c->device);
  WH_CUDA(cudaStreamSynchronize(c->stream));
  c->prof.on = on != 0 || wh::g_timeline != nullptr; c->prof.used = 0;
  if (on) c->prof.acc.clear();
  return WH_OK;
}
int wh_profile_count(whContext* c) { return c ? (int)c->prof.acc.size() : 0; }
int wh_profile_get(whContext* c, int i, const char** name, double* total_ms, uint64_t* count) {
  WH_REQUIRE(c && i >= 0 && i < (int)c->prof.acc.size(), "wh_profile_get: bad index");
  if (name) *name = c->prof.acc[i].name.c_str();
  if (total_ms) *total_ms = c->prof.acc[i].ms;
  if (count) *count = c->prof.acc[i].count;
  return WH_OK;
}

const char* wh_last_error(void) { return wh::g_err; }
const char* wh_version(void) { return "wormhole_b200 0.1 (sm_100a)"; }
uint64_t wh_kernel_launches(void) { return wh::g_launches.load(); }

int wh_create(int device, whContext** out) {
  WH_REQUIRE(out != nullptr, "wh_create: out is NULL");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    wh::set_error("wh_create: no CUDA device (%s); this library has no CPU path",
                  e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
    return WH_ERR_CUDA;
  }
  WH_REQUIRE(device >= 0 && device < n, "wh_create: device %d out of range [0,%d)", device, n);
  WH_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  WH_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) {
    wh::set_error("wh_create: device %d is sm_%d%d; the kernels are built for sm_100a only", device,
                  prop.major, prop.minor);
    return WH_ERR_UNSUPPORTED;
  }
  // WH_L2_FETCH=32|64|128: the L2's DRAM fetch granularity (a device-wide hint; the path's table accesses are random
  // 4- to 16-byte reads, of which a wider fetch only keeps one sector)
  if (const char* e = getenv("WH_L2_FETCH")) { const int g = atoi(e); if (g == 32 || g == 64 || g == 128) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)g); }
  whContext* c = new whContext();
  if (const char* tl = getenv("WH_TIMELINE")) {
    if (!wh::g_timeline) {
      const char* rk = getenv("RANK");
      wh::g_timeline = fopen((std::string(tl) + "." + (rk ? rk : "0")).c_str(), "w");
    }
    c->prof.on = wh::g_timeline != nullptr;
  }
  c->device = device;
  c->num_sms = prop.multiProcessorCount;
  c->l2_bytes = (size_t)prop.l2CacheSize;
  WH_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  c->own_stream = true;
  int r = c->pinned.reserve(1 << 16);
  if (r == WH_OK) r = c->small.reserve(1 << 16);
  if (r != WH_OK) { delete c; return r; }
  *out = c;
  return WH_OK;
}

int wh_destroy(whContext* c) {
  if (!c) return WH_OK;
  wh::DeviceGuard g(c->device);
  cudaStreamSynchronize(c->stream);
  if (c->side_ev) cudaEventDestroy(c->side_ev);
  c->directory.release(); c->scan_tmp.release(); c->scan_state.release(); c->acc.release(); c->auc.release(); c->small.release();
  for (auto& b : c->sort_tmp) b.release();
  for (auto& b : c->stage) b.release();
  c->pinned.release();
  for (auto e : c->prof.ev) cudaEventDestroy(e);
  if (c->own_stream) cudaStreamDestroy(c->stream);
  delete c;
  return WH_OK;
}

int wh_set_stream(whContext* c, void* s) {
  WH_REQUIRE(c != nullptr, "wh_set_stream: ctx is NULL");
  wh::DeviceGuard g(c->device);
  WH_CUDA(cudaStreamSynchronize(c->stream));
  if (c->own_stream) cudaStreamDestroy(c->stream);
  c->stream = (cudaStream_t)s;
  c->own_stream = false;
  return WH_OK;
}

int wh_sync(whContext* c) {
  WH_REQUIRE(c != nullptr, "wh_sync: ctx is NULL");
  wh::DeviceGuard g(c->device);
  WH_CUDA(cudaStreamSynchronize(c->stream));
  return WH_OK;
}

uint64_t wh_reverse_bytes(uint64_t x) { return wh::reverse_nibbles(x); }

}  // extern "C"
