// dgs_l2norm_noalloc: the two-pass global L2 norm for callers that cannot allocate device memory.
//
// Who needs this.  dgs_l2norm() asks its caller for a scratch buffer (one fp32 partial per CTA of pass 1).
// Some callers have no way to provide one at the moment they need the norm:
//   * a plugin host that binds this library over cgo / JNI / ctypes and owns no device allocator -- it is
//     handed the gradient buffer and an output slot by the training process and nothing else;
//   * a trainer whose caching allocator has already reserved (nearly) all of HBM, where a raw cudaMalloc
//     for a few KB of scratch can fail with out-of-memory although the framework itself still has room;
//   * code that runs under a "no allocation after start-up" rule (latency-sensitive step loops, memory
//     accounting that must not move once training has begun).
// For them the scratch lives inside the library, in a fixed-size __device__ array that is part of the
// loaded module.  Every GPU that loads libdgs.so holds DGS_NOALLOC_MAX_PARTIALS * 4 bytes (16 KiB) for it
// until the process exits, whether or not this entry point is ever called.
//
// What the caller must know.
//   * Limit: n <= DGS_NOALLOC_MAX_ELEMS (2^26 elements); larger inputs are refused with DGS_EINVAL, they
//     are never truncated.  Use dgs_l2norm() with your own scratch above that.
//   * The first call on each device looks up the device address of the array (cudaGetSymbolAddress, with
//     that device current) and caches it.  Make that first call during set-up, before any stream capture
//     is started on that device; later calls only launch kernels.
//   * No cudaMalloc, cudaFree or other allocating routine is called here or below, on any call.
//   * There is ONE array per device, shared by every call and every stream.  Two calls whose kernels
//     overlap on the GPU (different streams, no ordering between them) race on it and may both return a
//     wrong norm; order such calls yourself.  Calls on one stream are safe.
//   * The array keeps whatever the previous call left in it.  Nothing in it has to start at zero: pass 1
//     overwrites every slot that pass 2 of the same call reads, so no clearing is done or needed.
//
// There is no counterpart of this entry point in the reference.
#include "dgs_common.cuh"

#include <mutex>

namespace {

// Host code must never pass this name to a kernel: on the host it is only a handle for the symbol, not a
// device address.  dgs_l2norm_noalloc() resolves the address once per device (below) and passes that.
static __device__ __align__(256) float g_noalloc_partials[DGS_NOALLOC_MAX_PARTIALS];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
float* g_partials_addr[kMaxDevices] = {nullptr};

// Device address of g_noalloc_partials on the CURRENT device; resolved on the first call per device.
int scratch_for_current_device(float** out) {
  int dev = -1;
  DGS_CUDA(cudaGetDevice(&dev));
  DGS_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "dgs_l2norm_noalloc: device index %d is outside [0, %d)", dev,
                kMaxDevices);
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (g_partials_addr[dev] == nullptr) {
    float* addr = nullptr;
    DGS_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&addr), g_noalloc_partials));
    DGS_CHECK_ARG(addr != nullptr, "dgs_l2norm_noalloc: cudaGetSymbolAddress returned a null address");
    g_partials_addr[dev] = addr;
  }
  *out = g_partials_addr[dev];
  return DGS_OK;
}

}  // namespace

extern "C" int dgs_l2norm_noalloc(const void* x, int dtype, int64_t n, float* out_norm, dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(n >= 0, "dgs_l2norm_noalloc: n < 0");
  DGS_CHECK_ARG(n <= (int64_t)DGS_NOALLOC_MAX_ELEMS,
                "dgs_l2norm_noalloc: n = %lld exceeds DGS_NOALLOC_MAX_ELEMS = %lld (the built-in scratch holds %d "
                "partials); call dgs_l2norm() with caller-provided scratch for larger buffers",
                (long long)n, (long long)DGS_NOALLOC_MAX_ELEMS, DGS_NOALLOC_MAX_PARTIALS);
  DGS_CHECK_ARG(out_norm != nullptr, "dgs_l2norm_noalloc: out_norm is NULL");
  const int grid = dgs_l2norm_grid(n);
  DGS_CHECK_ARG(grid <= DGS_NOALLOC_MAX_PARTIALS, "dgs_l2norm_noalloc: %d partials do not fit the built-in scratch", grid);
  float* scratch = nullptr;
  if (int rc = scratch_for_current_device(&scratch)) return rc;
  // same two kernels as dgs_l2norm(), with the library-owned scratch
  if (int rc = dgs_l2norm_partial(x, dtype, n, scratch, stream)) return rc;
  return dgs_l2norm_final(scratch, grid, out_norm, stream);
}
