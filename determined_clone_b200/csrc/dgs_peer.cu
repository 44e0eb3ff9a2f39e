// Peer-memory exchange over NVLink 5 / NVSwitch: the bucket "allreduce" of the data-parallel path done by this
// library's own kernels on peer-mapped buffers instead of ncclAllReduce.
//
// Takes over DDP's bucket allreduce on the boundary backward pass (reference
// harness/determined/pytorch/_pytorch_context.py:295-296 wraps the model in DistributedDataParallel, :813-816 is the
// backward pass whose hooks all-reduce the buckets) for up to DGS_PEER_MAX ranks on one NVSwitch domain:
//
//   backward, per bucket   dgs_flatten_push      grads (+ fp32 accumulator) * 1/(N*A) -> cast -> written straight into
//                                                the OWNER rank's wire buffer (row = source rank): flatten, scale, cast
//                                                and the reduce-scatter transfer are one kernel, posted stores over NVLink
//   step_optimizer         dgs_peer_barrier      every rank's pushes have landed
//                          dgs_peer_reduce       owner sums its N rows in rank order (deterministic; fp32 accumulate,
//                                                one rounding) and emits the per-CTA sums of squares of the L2 norm
//                          dgs_peer_norm_exchange  8-byte all-to-all of the ranks' sums of squares + barrier -> norm
//                          dgs_step_*_push       (dgs_step.cu) clip + optimizer step on the owner's 1/N slice, new
//                                                parameters stored to every rank's parameter buffer (the all-gather)
//                          dgs_peer_barrier      every rank's parameter stores have landed
//
// Every rank maps the others' buffers with CUDA IPC (cudaIpcGetMemHandle / cudaIpcOpenMemHandle); the 64-byte handles
// travel through the caller's control plane.  Cross-GPU ordering uses flags in a small peer-mapped sync block per rank:
// st.release.sys / ld.acquire.sys with a monotonically increasing epoch per slot, so nothing is ever reset.  A wait
// that exceeds its timeout does not trap (that would be a GPU fault): it records an error code in a host-mapped word,
// stops waiting, and every later wait returns at once, so the stream drains and the host can raise.
#include "dgs_common.cuh"

#include <cuda.h>

#include <cstring>
#include <map>
#include <mutex>
#include <string>

namespace dgs {
namespace {

constexpr int kThreads = 256;

// ---- IPC bookkeeping ----------------------------------------------------------------------------------------------
std::mutex g_ipc_mutex;
struct Opened {
  void* base;
  int refs;
};
std::map<std::string, Opened> g_opened;  // handle bytes -> mapped base pointer (one open per allocation per process)
int* g_err_host = nullptr;              // host-mapped error word, shared by every peer group of the process
int* g_err_dev = nullptr;

int ensure_error_word() {
  if (g_err_host != nullptr) return DGS_OK;
  DGS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&g_err_host), sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable));
  *g_err_host = 0;
  DGS_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&g_err_dev), g_err_host, 0));
  return DGS_OK;
}

// ---- device-side flag protocol ----------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

struct SyncPtrs {
  unsigned int* flags[DGS_PEER_MAX];  // flags[r] = rank r's sync block (peer-mapped; [me] is local)
};

// sync block layout (uint32 words): flag[slot][src] at slot * DGS_PEER_MAX + src, then doubles for the norm exchange
constexpr int kFlagWords = DGS_PEER_SLOTS * DGS_PEER_MAX;
__device__ __forceinline__ double* sumsq_slots(unsigned int* block, unsigned int epoch) {
  // double-buffered by epoch parity: a fast rank may write epoch e+1 while a slow one still reads epoch e
  return reinterpret_cast<double*>(block + kFlagWords) + (epoch & 1u) * DGS_PEER_MAX;
}

// thread `t` < nranks: tell rank t that this rank reached `epoch` on `slot`, then wait for rank t's flag here.
// Epoch comparison is wrap-safe (signed difference).
__device__ __forceinline__ void signal_and_wait(const SyncPtrs& S, int slot, unsigned int epoch, int nranks, int rank,
                                                unsigned long long timeout_ns, int* err) {
  const int t = threadIdx.x;
  if (t < nranks) {
    __threadfence_system();  // everything this rank wrote before (earlier kernels included) is visible system-wide first
    st_release_sys(S.flags[t] + slot * DGS_PEER_MAX + rank, epoch);
    const unsigned int* mine = S.flags[rank] + slot * DGS_PEER_MAX + t;
    unsigned long long t0 = 0;
    unsigned int spins = 0;
    while ((int)(ld_acquire_sys(mine) - epoch) < 0) {
      // the slow path only: the error word lives in host memory (a PCIe read), so it is looked at once the wait has
      // lasted a while -- a wait that follows an earlier timeout then ends after ~1 ms instead of a full timeout
      if ((++spins & 1023u) == 0) {
        const unsigned long long now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        if (*reinterpret_cast<volatile int*>(err) != 0) break;
        if (now - t0 > timeout_ns) {
          *reinterpret_cast<volatile int*>(err) = 1000 + slot * 16 + t;  // which slot, which peer never arrived
          __threadfence_system();
          break;
        }
      }
      __nanosleep(32);
    }
  }
  __syncthreads();
}

__global__ void barrier_kernel(const __grid_constant__ SyncPtrs S, int slot, unsigned int epoch, int nranks, int rank,
                               unsigned long long timeout_ns, int* err) {
  signal_and_wait(S, slot, epoch, nranks, rank, timeout_ns, err);
}

// One CTA: sum this rank's per-CTA partial sums of squares in double, hand the result to every rank, wait for
// everyone's, add them in rank order (the same order on every rank: identical norms bit for bit), take the root.
__global__ void __launch_bounds__(kThreads)
norm_exchange_kernel(const __grid_constant__ SyncPtrs S, const float* __restrict__ partials, int n_partials, int slot,
                     unsigned int epoch, int nranks, int rank, unsigned long long timeout_ns, int* err,
                     float* __restrict__ out_norm) {
  __shared__ double red[32];
  double s = 0.0;
  for (int i = threadIdx.x; i < n_partials; i += kThreads) s += (double)partials[i];
  s = block_sum(s, red);
  __shared__ double mine;
  if (threadIdx.x == 0) mine = s;
  __syncthreads();
  if (threadIdx.x < nranks) sumsq_slots(S.flags[threadIdx.x], epoch)[rank] = mine;  // plain store, released by the flag below
  signal_and_wait(S, slot, epoch, nranks, rank, timeout_ns, err);
  if (threadIdx.x == 0) {
    const double* slots = sumsq_slots(S.flags[rank], epoch);
    double tot = 0.0;
    for (int r = 0; r < nranks; ++r) tot += reinterpret_cast<const volatile double*>(slots)[r];
    out_norm[0] = (float)sqrt(tot);
  }
}

// Owner-side reduction: rows r = 0..N-1 of this bucket in MY wire hold rank r's contribution to my slice.
// wire[begin + rank*span + k] = sum_r wire[begin + r*span + k] (fp32 accumulate in rank order), plus the sum of
// squares of the result for the global norm.  Reads N*span elements once, writes span: local HBM only.
template <typename W>
__global__ void __launch_bounds__(kThreads)
peer_reduce_kernel(W* __restrict__ bucket, long long span, int nranks, int rank, float* __restrict__ partials) {
  constexpr int VEC = is16<W>::value ? 8 : 4;
  __shared__ float red[32];
  const long long nvec = span / VEC;  // span is a multiple of 8 elements
  const long long stride = (long long)gridDim.x * kThreads;
  float ss = 0.f;
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < nvec; i += stride) {
    float acc[VEC], v[DGS_PEER_MAX][VEC];
#pragma unroll
    for (int r = 0; r < DGS_PEER_MAX; ++r)
      if (r < nranks) load_vec<W, VEC, Cache::Stream>(bucket + r * span + i * VEC, v[r]);
#pragma unroll
    for (int k = 0; k < VEC; ++k) acc[k] = v[0][k];
#pragma unroll
    for (int r = 1; r < DGS_PEER_MAX; ++r)
      if (r < nranks) {
#pragma unroll
        for (int k = 0; k < VEC; ++k) acc[k] += v[r][k];
      }
    store_vec<W, VEC>(bucket + (long long)rank * span + i * VEC, acc);
    // the norm is that of the values as stored (rounded to the wire dtype), as clip_grad_norm_ would see them
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      const float q = to_f32(from_f32<W>(acc[k]));
      ss = fmaf(q, q, ss);
    }
  }
  const float tot = block_sum(ss, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = tot;
}

// The same reduction over every bucket of a flat buffer in one launch: tiles of all buckets' slices are dealt to a
// persistent grid, each CTA leaves ONE partial (so the norm needs gridDim.x slots however many buckets there are).
struct ReduceTable {
  long long begin[DGS_PEER_MAX_BUCKETS];       // element offset of the bucket in the wire buffer
  long long span[DGS_PEER_MAX_BUCKETS];        // elements per rank
  long long tile_start[DGS_PEER_MAX_BUCKETS];  // first tile of the bucket's slice
  int n;
  long long tiles;
};

template <typename W>
__global__ void __launch_bounds__(kThreads)
peer_reduce_multi_kernel(W* __restrict__ wire, const __grid_constant__ ReduceTable T, int nranks, int rank,
                         float* __restrict__ partials, int capacity) {
  constexpr int VEC = is16<W>::value ? 8 : 4;
  constexpr int kU = 2;
  constexpr long long TILE = (long long)kThreads * VEC * kU;
  __shared__ float red[32];
  float ss = 0.f;
  for (long long t = blockIdx.x; t < T.tiles; t += gridDim.x) {
    int lo = 0, hi = T.n - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (T.tile_start[mid] <= t) lo = mid; else hi = mid - 1;
    }
    W* __restrict__ bucket = wire + T.begin[lo];
    const long long span = T.span[lo];
    const long long base = (t - T.tile_start[lo]) * TILE;
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = base + (long long)(u * kThreads + threadIdx.x) * VEC;
      if (i < span) {  // span is a multiple of 8: a vector is inside or outside as a whole
        float acc[VEC], v[DGS_PEER_MAX][VEC];
#pragma unroll
        for (int r = 0; r < DGS_PEER_MAX; ++r)
          if (r < nranks) load_vec<W, VEC, Cache::Stream>(bucket + r * span + i, v[r]);
#pragma unroll
        for (int k = 0; k < VEC; ++k) acc[k] = v[0][k];
#pragma unroll
        for (int r = 1; r < DGS_PEER_MAX; ++r)
          if (r < nranks) {
#pragma unroll
            for (int k = 0; k < VEC; ++k) acc[k] += v[r][k];
          }
        store_vec<W, VEC>(bucket + (long long)rank * span + i, acc);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
          const float q = to_f32(from_f32<W>(acc[k]));
          ss = fmaf(q, q, ss);
        }
      }
    }
  }
  const float tot = block_sum(ss, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = tot;
  // slots this launch does not use must read as zero in the final sum
  if (blockIdx.x == 0)
    for (int i = gridDim.x + threadIdx.x; i < capacity; i += kThreads) partials[i] = 0.f;
}

// All-gather by stores: pieces of this rank's buffer (its slices of every bucket) copied to the same offsets of every
// other rank's buffer.  One CTA moves one 8 KB tile to ONE destination; consecutive CTAs go to different destinations
// (rotated by rank), so that the CTAs in flight at any moment load every link evenly.
struct ScatterTable {
  long long off[DGS_PEER_MAX_BUCKETS];         // byte offset of the piece (multiple of 16)
  long long bytes[DGS_PEER_MAX_BUCKETS];       // multiple of 16
  long long tile_start[DGS_PEER_MAX_BUCKETS];
  int n;
  long long tiles;
};
struct PeerBases {
  char* p[DGS_PEER_MAX];
};
constexpr long long kScatterTile = (long long)kThreads * 16 * 2;

__global__ void __launch_bounds__(kThreads)
peer_scatter_kernel(const __grid_constant__ PeerBases B, const __grid_constant__ ScatterTable T, int nranks, int rank) {
  const long long n_visit = T.tiles * (nranks - 1);
  for (long long i = blockIdx.x; i < n_visit; i += gridDim.x) {
    const long long t = i / (nranks - 1);
    const int dest = (int)((i % (nranks - 1) + rank + 1) % nranks);
    int lo = 0, hi = T.n - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (T.tile_start[mid] <= t) lo = mid; else hi = mid - 1;
    }
    const long long base = T.off[lo] + (t - T.tile_start[lo]) * kScatterTile;
    const long long end = T.off[lo] + T.bytes[lo];
    const char* __restrict__ src = B.p[rank];
    char* __restrict__ dst = B.p[dest];
    uint4 v[2];
    long long at[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      at[u] = base + ((long long)u * kThreads + threadIdx.x) * 16;
      if (at[u] < end) v[u] = *reinterpret_cast<const uint4*>(src + at[u]);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u)
      if (at[u] < end) *reinterpret_cast<uint4*>(dst + at[u]) = v[u];
  }
}

int check_group(const char* who, int nranks, int rank) {
  DGS_CHECK_ARG(nranks >= 2 && nranks <= DGS_PEER_MAX && rank >= 0 && rank < nranks, "%s: rank %d of %d (at most %d ranks)",
                who, rank, nranks, DGS_PEER_MAX);
  return DGS_OK;
}

int fill_sync(const char* who, void* const* sync_blocks, int nranks, SyncPtrs* S) {
  DGS_CHECK_ARG(sync_blocks != nullptr, "%s: sync_blocks is NULL", who);
  for (int r = 0; r < DGS_PEER_MAX; ++r) S->flags[r] = nullptr;
  for (int r = 0; r < nranks; ++r) {
    DGS_CHECK_ARG(sync_blocks[r] != nullptr && (reinterpret_cast<uintptr_t>(sync_blocks[r]) & 15u) == 0,
                  "%s: sync block of rank %d is NULL or misaligned", who, r);
    S->flags[r] = static_cast<unsigned int*>(sync_blocks[r]);
  }
  return DGS_OK;
}

unsigned long long timeout_ns(double seconds) {
  if (!(seconds > 0)) seconds = 120.0;
  return (unsigned long long)(seconds * 1e9);
}

}  // namespace
}  // namespace dgs

extern "C" int dgs_peer_export(const void* ptr, void* handle64_out, int64_t* offset_out) {
  using namespace dgs;
  static_assert(sizeof(cudaIpcMemHandle_t) == DGS_PEER_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
  DGS_CHECK_ARG(ptr && handle64_out && offset_out, "dgs_peer_export: NULL argument");
  // a handle names a whole allocation: find the allocation `ptr` lies in (driver API through the runtime's loader, so
  // that the library needs no link-time libcuda and still loads on machines without a driver)
  typedef CUresult (*GetRange)(CUdeviceptr*, size_t*, CUdeviceptr);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  DGS_CUDA(cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &q));
  DGS_CHECK_ARG(fn != nullptr && q == cudaDriverEntryPointSuccess, "dgs_peer_export: cuMemGetAddressRange is unavailable");
  CUdeviceptr base = 0;
  size_t size = 0;
  const CUresult rc = reinterpret_cast<GetRange>(fn)(&base, &size, reinterpret_cast<CUdeviceptr>(ptr));
  if (rc != CUDA_SUCCESS) return fail(DGS_ECUDA, "dgs_peer_export: cuMemGetAddressRange failed (%d)", (int)rc);
  cudaIpcMemHandle_t h;
  DGS_CUDA(cudaIpcGetMemHandle(&h, reinterpret_cast<void*>(base)));
  memcpy(handle64_out, &h, sizeof(h));
  *offset_out = (int64_t)(reinterpret_cast<CUdeviceptr>(ptr) - base);
  return DGS_OK;
}

extern "C" int dgs_peer_open(const void* handle64, int64_t offset, void** out_ptr) {
  using namespace dgs;
  DGS_CHECK_ARG(handle64 && out_ptr && offset >= 0, "dgs_peer_open: bad argument");
  std::lock_guard<std::mutex> lock(g_ipc_mutex);
  const std::string key(static_cast<const char*>(handle64), DGS_PEER_HANDLE_BYTES);
  auto it = g_opened.find(key);
  void* base = nullptr;
  if (it != g_opened.end()) {
    base = it->second.base;
    ++it->second.refs;
  } else {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    DGS_CUDA(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
    g_opened.emplace(key, Opened{base, 1});
  }
  *out_ptr = static_cast<char*>(base) + offset;
  return DGS_OK;
}

extern "C" int dgs_peer_close(const void* handle64) {
  using namespace dgs;
  DGS_CHECK_ARG(handle64 != nullptr, "dgs_peer_close: NULL argument");
  std::lock_guard<std::mutex> lock(g_ipc_mutex);
  auto it = g_opened.find(std::string(static_cast<const char*>(handle64), DGS_PEER_HANDLE_BYTES));
  DGS_CHECK_ARG(it != g_opened.end(), "dgs_peer_close: this handle is not open");
  if (--it->second.refs == 0) {
    const cudaError_t e = cudaIpcCloseMemHandle(it->second.base);
    g_opened.erase(it);
    if (e != cudaSuccess) return fail(DGS_ECUDA, "cudaIpcCloseMemHandle failed: %s", cudaGetErrorString(e));
  }
  return DGS_OK;
}

extern "C" int dgs_peer_close_all(void) {
  using namespace dgs;
  std::lock_guard<std::mutex> lock(g_ipc_mutex);
  for (auto& kv : g_opened) (void)cudaIpcCloseMemHandle(kv.second.base);
  g_opened.clear();
  (void)cudaGetLastError();
  return DGS_OK;
}

extern "C" int dgs_peer_error(int* code_out) {
  using namespace dgs;
  DGS_CHECK_ARG(code_out != nullptr, "dgs_peer_error: NULL argument");
  *code_out = (g_err_host != nullptr) ? *reinterpret_cast<volatile int*>(g_err_host) : 0;
  return DGS_OK;
}

extern "C" int dgs_peer_error_clear(void) {
  if (dgs::g_err_host != nullptr) *reinterpret_cast<volatile int*>(dgs::g_err_host) = 0;
  return DGS_OK;
}

extern "C" int dgs_peer_barrier(void* const* sync_blocks, int nranks, int rank, int slot, uint32_t epoch,
                                double timeout_s, dgs_stream_t stream) {
  using namespace dgs;
  if (int rc = check_group("dgs_peer_barrier", nranks, rank)) return rc;
  DGS_CHECK_ARG(slot >= 0 && slot < DGS_PEER_SLOTS, "dgs_peer_barrier: slot %d outside [0, %d)", slot, DGS_PEER_SLOTS);
  SyncPtrs S;
  if (int rc = fill_sync("dgs_peer_barrier", sync_blocks, nranks, &S)) return rc;
  if (int rc = ensure_error_word()) return rc;
  barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(S, slot, epoch, nranks, rank, timeout_ns(timeout_s), g_err_dev);
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_peer_norm_exchange(void* const* sync_blocks, int nranks, int rank, int slot, uint32_t epoch,
                                      const float* partials, int n_partials, float* out_norm, double timeout_s,
                                      dgs_stream_t stream) {
  using namespace dgs;
  if (int rc = check_group("dgs_peer_norm_exchange", nranks, rank)) return rc;
  DGS_CHECK_ARG(slot >= 0 && slot < DGS_PEER_SLOTS, "dgs_peer_norm_exchange: slot %d outside [0, %d)", slot, DGS_PEER_SLOTS);
  DGS_CHECK_ARG(partials != nullptr && out_norm != nullptr && n_partials >= 0, "dgs_peer_norm_exchange: bad buffers");
  SyncPtrs S;
  if (int rc = fill_sync("dgs_peer_norm_exchange", sync_blocks, nranks, &S)) return rc;
  if (int rc = ensure_error_word()) return rc;
  norm_exchange_kernel<<<1, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(S, partials, n_partials, slot, epoch, nranks, rank,
                                                                                timeout_ns(timeout_s), g_err_dev, out_norm);
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_peer_reduce(void* bucket, int dtype, int64_t span, int nranks, int rank, float* partials,
                               dgs_stream_t stream) {
  using namespace dgs;
  if (int rc = check_group("dgs_peer_reduce", nranks, rank)) return rc;
  DGS_CHECK_ARG(span >= 0 && (span & 7) == 0, "dgs_peer_reduce: span %lld is not a multiple of 8 elements", (long long)span);
  DGS_CHECK_ARG(partials != nullptr, "dgs_peer_reduce: partials is NULL");
  DGS_CHECK_ARG(span == 0 || (bucket != nullptr && aligned16(bucket)), "dgs_peer_reduce: bucket is NULL or misaligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = dgs_l2norm_grid(span);
  switch (dtype) {
    case DGS_F32: peer_reduce_kernel<float><<<grid, kThreads, 0, st>>>((float*)bucket, span, nranks, rank, partials); break;
    case DGS_BF16: peer_reduce_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>((__nv_bfloat16*)bucket, span, nranks, rank, partials); break;
    case DGS_F16: peer_reduce_kernel<__half><<<grid, kThreads, 0, st>>>((__half*)bucket, span, nranks, rank, partials); break;
    default: return fail(DGS_EINVAL, "dgs_peer_reduce: bad dtype %d", dtype);
  }
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_peer_reduce_multi(void* wire, int dtype, const int64_t* bucket_begin, const int64_t* bucket_span,
                                     int n_buckets, int nranks, int rank, float* partials, int partials_capacity,
                                     dgs_stream_t stream) {
  using namespace dgs;
  if (int rc = check_group("dgs_peer_reduce_multi", nranks, rank)) return rc;
  DGS_CHECK_ARG(n_buckets >= 0 && n_buckets <= DGS_PEER_MAX_BUCKETS, "dgs_peer_reduce_multi: n_buckets %d outside [0, %d]", n_buckets,
                DGS_PEER_MAX_BUCKETS);
  DGS_CHECK_ARG(partials != nullptr && partials_capacity >= 1, "dgs_peer_reduce_multi: partials scratch is missing");
  DGS_CHECK_ARG(n_buckets == 0 || (wire && aligned16(wire) && bucket_begin && bucket_span), "dgs_peer_reduce_multi: NULL or misaligned buffer");
  const int vec = (dtype == DGS_F32) ? 4 : 8;
  const long long tile = (long long)kThreads * vec * 2;
  ReduceTable T;
  T.n = 0;
  T.tiles = 0;
  for (int b = 0; b < n_buckets; ++b) {
    DGS_CHECK_ARG(bucket_begin[b] >= 0 && (bucket_begin[b] & 7) == 0 && bucket_span[b] >= 0 && (bucket_span[b] & 7) == 0,
                  "dgs_peer_reduce_multi: bucket %d: begin %lld / span %lld must be multiples of 8 elements", b,
                  (long long)bucket_begin[b], (long long)bucket_span[b]);
    if (bucket_span[b] == 0) continue;
    T.begin[T.n] = bucket_begin[b];
    T.span[T.n] = bucket_span[b];
    T.tile_start[T.n] = T.tiles;
    T.tiles += (bucket_span[b] + tile - 1) / tile;
    ++T.n;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  long long grid = (long long)sm_count() * 8;
  if (grid > T.tiles) grid = T.tiles;
  if (grid > partials_capacity) grid = partials_capacity;
  if (grid < 1) grid = 1;
  switch (dtype) {
    case DGS_F32: peer_reduce_multi_kernel<float><<<(int)grid, kThreads, 0, st>>>((float*)wire, T, nranks, rank, partials, partials_capacity); break;
    case DGS_BF16: peer_reduce_multi_kernel<__nv_bfloat16><<<(int)grid, kThreads, 0, st>>>((__nv_bfloat16*)wire, T, nranks, rank, partials, partials_capacity); break;
    case DGS_F16: peer_reduce_multi_kernel<__half><<<(int)grid, kThreads, 0, st>>>((__half*)wire, T, nranks, rank, partials, partials_capacity); break;
    default: return fail(DGS_EINVAL, "dgs_peer_reduce_multi: bad dtype %d", dtype);
  }
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_peer_scatter(void* const* bases, const int64_t* piece_off, const int64_t* piece_bytes, int n_pieces,
                                int nranks, int rank, dgs_stream_t stream) {
  using namespace dgs;
  if (int rc = check_group("dgs_peer_scatter", nranks, rank)) return rc;
  DGS_CHECK_ARG(n_pieces >= 0 && n_pieces <= DGS_PEER_MAX_BUCKETS, "dgs_peer_scatter: n_pieces %d outside [0, %d]", n_pieces,
                DGS_PEER_MAX_BUCKETS);
  DGS_CHECK_ARG(bases != nullptr && (n_pieces == 0 || (piece_off && piece_bytes)), "dgs_peer_scatter: NULL argument");
  PeerBases B;
  for (int r = 0; r < DGS_PEER_MAX; ++r) B.p[r] = nullptr;
  for (int r = 0; r < nranks; ++r) {
    DGS_CHECK_ARG(bases[r] != nullptr && aligned16(bases[r]), "dgs_peer_scatter: buffer of rank %d is NULL or misaligned", r);
    B.p[r] = static_cast<char*>(bases[r]);
  }
  ScatterTable T;
  T.n = 0;
  T.tiles = 0;
  for (int i = 0; i < n_pieces; ++i) {
    DGS_CHECK_ARG(piece_off[i] >= 0 && (piece_off[i] & 15) == 0 && piece_bytes[i] >= 0 && (piece_bytes[i] & 15) == 0,
                  "dgs_peer_scatter: piece %d: offset %lld / size %lld must be multiples of 16 bytes", i, (long long)piece_off[i],
                  (long long)piece_bytes[i]);
    if (piece_bytes[i] == 0) continue;
    T.off[T.n] = piece_off[i];
    T.bytes[T.n] = piece_bytes[i];
    T.tile_start[T.n] = T.tiles;
    T.tiles += (piece_bytes[i] + kScatterTile - 1) / kScatterTile;
    ++T.n;
  }
  if (T.tiles == 0) return DGS_OK;
  const long long visit = T.tiles * (nranks - 1);
  const long long cap = (long long)sm_count() * grid_mult();
  peer_scatter_kernel<<<(int)(visit < cap ? visit : cap), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(B, T, nranks, rank);
  DGS_LAUNCHED();
  return DGS_OK;
}
