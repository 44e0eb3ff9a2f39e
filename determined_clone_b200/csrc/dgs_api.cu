// Error reporting, launch accounting and device queries shared by every entry point of libdgs.so.
#include "dgs_common.cuh"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <utility>
#include <vector>

namespace dgs {

std::atomic<int64_t> g_launches{0};

namespace {
thread_local char t_error[512] = "";
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_error, sizeof(t_error), fmt, ap);
  va_end(ap);
  return code;
}

int sm_count() {
  constexpr int kMaxDev = 64;
  static std::atomic<int> cache[kMaxDev];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDev) {
    (void)cudaGetLastError();
    return 148;  // B200; only reached for host-side sizing queries on a machine without a GPU
  }
  int v = cache[dev].load(std::memory_order_relaxed);
  if (v == 0) {
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) {
      (void)cudaGetLastError();
      v = 148;
    }
    cache[dev].store(v, std::memory_order_relaxed);
  }
  return v;
}

int grid_mult() {
  static const int v = [] {
    const char* e = getenv("DGS_GRID_MULT");
    int m = e ? atoi(e) : 0;
    // Measured on B200 (gpurun_out/sweep_grid.txt): one tile per CTA beats a persistent grid-stride loop
    // by 5-15% on mixed read/write streams (CTAs stay in address order, so DRAM pages stay open).
    return m > 0 ? m : 4096;
  }();
  return v;
}

// ---- event timing of the step kernels -----------------------------------------------------------------
namespace {
constexpr int kMaxTimed = 1024;
std::atomic<bool> g_timing{false};
std::mutex g_timer_mutex;
std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_pool;  // created on first use, reused for ever
int g_used = 0;
bool g_open = false;
}  // namespace

void timer_begin(cudaStream_t st) {
  if (!g_timing.load(std::memory_order_relaxed)) return;
  std::lock_guard<std::mutex> lock(g_timer_mutex);
  if (g_used >= kMaxTimed) return;
  if (g_used >= (int)g_pool.size()) {
    cudaEvent_t a, b;
    if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) {
      (void)cudaGetLastError();
      return;
    }
    g_pool.emplace_back(a, b);
  }
  g_open = cudaEventRecord(g_pool[g_used].first, st) == cudaSuccess;
}

void timer_end(cudaStream_t st) {
  if (!g_timing.load(std::memory_order_relaxed)) return;
  std::lock_guard<std::mutex> lock(g_timer_mutex);
  if (!g_open) return;
  g_open = false;
  if (cudaEventRecord(g_pool[g_used].second, st) == cudaSuccess) ++g_used;
}

}  // namespace dgs

extern "C" int dgs_timer_enable(int enabled) {
  dgs::g_timing.store(enabled != 0);
  return DGS_OK;
}

extern "C" int dgs_timer_read(float* us_out, int capacity, int* n_out) {
  using namespace dgs;
  DGS_CHECK_ARG(n_out != nullptr && capacity >= 0 && (capacity == 0 || us_out != nullptr), "dgs_timer_read: bad arguments");
  std::lock_guard<std::mutex> lock(g_timer_mutex);
  int n = 0;
  if (g_used > 0) {
    DGS_CUDA(cudaEventSynchronize(g_pool[g_used - 1].second));
    for (int i = 0; i < g_used && n < capacity; ++i) {
      float ms = 0.f;
      DGS_CUDA(cudaEventElapsedTime(&ms, g_pool[i].first, g_pool[i].second));
      us_out[n++] = ms * 1000.f;
    }
  }
  g_used = 0;
  *n_out = n;
  return DGS_OK;
}

extern "C" int dgs_version(void) { return DGS_VERSION; }
extern "C" const char* dgs_last_error(void) { return dgs::t_error; }
extern "C" int64_t dgs_launch_count(void) { return dgs::g_launches.load(); }
