// Error reporting, launch accounting and device queries shared by every entry point of libdgs.so.
#include "dgs_common.cuh"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>

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

}  // namespace dgs

extern "C" int dgs_version(void) { return DGS_VERSION; }
extern "C" const char* dgs_last_error(void) { return dgs::t_error; }
extern "C" int64_t dgs_launch_count(void) { return dgs::g_launches.load(); }
