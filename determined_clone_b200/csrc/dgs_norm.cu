// Two-pass global L2 norm over a flat gradient buffer, and the unfused clip tail.
//
// Takes over torch.nn.utils.clip_grad_norm_ as called through the user's clip_grads callable at
// reference harness/determined/pytorch/_pytorch_context.py:902 (torch/nn/utils/clip_grad.py:
// _get_total_norm = norm of per-tensor norms == sqrt of the flat sum of squares).
//
// pass 1 reads the buffer once (HBM-bound): per-thread fp32 partial sums over 4 independent 16 B
// loads, warp shuffle, shared-memory block reduction, one fp32 partial per CTA.  pass 2 is one CTA
// summing <= DGS_NORM_MAX_PARTIALS partials in double.  Nothing returns to the host.
#include "dgs_common.cuh"

#include <cstdlib>

namespace dgs {
namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;

template <typename T>
__global__ void __launch_bounds__(kThreads)
sumsq_kernel(const T* __restrict__ x, long long n, float* __restrict__ partials) {
  constexpr int VEC = is16<T>::value ? 8 : 4;
  __shared__ float red[32];
  const long long nvec = n / VEC;
  const long long stride = (long long)gridDim.x * kThreads;
  long long i = (long long)blockIdx.x * kThreads + threadIdx.x;
  float s[kUnroll] = {0.f, 0.f, 0.f, 0.f};
  // main loop: kUnroll independent vector loads in flight per thread
  for (; i + (kUnroll - 1) * stride < nvec; i += kUnroll * stride) {
    float v[kUnroll][VEC];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) load_vec<T, VEC>(x + (i + u * stride) * VEC, v[u]);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u)
#pragma unroll
      for (int k = 0; k < VEC; ++k) s[u] = fmaf(v[u][k], v[u][k], s[u]);
  }
  for (; i < nvec; i += stride) {
    float v[VEC];
    load_vec<T, VEC>(x + i * VEC, v);
#pragma unroll
    for (int k = 0; k < VEC; ++k) s[0] = fmaf(v[k], v[k], s[0]);
  }
  if (blockIdx.x == 0 && threadIdx.x < (int)(n - nvec * VEC)) {
    const float t = to_f32(x[nvec * VEC + threadIdx.x]);
    s[1] = fmaf(t, t, s[1]);
  }
  float tot = block_sum((s[0] + s[1]) + (s[2] + s[3]), red);
  if (threadIdx.x == 0) partials[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(kThreads)
norm_final_kernel(const float* __restrict__ partials, int n, float* __restrict__ out) {
  __shared__ double red[32];
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += kThreads) s += (double)partials[i];
  s = block_sum(s, red);
  if (threadIdx.x == 0) out[0] = (float)sqrt(s);
}

__global__ void __launch_bounds__(kThreads)
norm_final_sumsq_kernel(const float* __restrict__ partials, int n, float* __restrict__ out) {
  __shared__ double red[32];
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += kThreads) s += (double)partials[i];
  s = block_sum(s, red);
  if (threadIdx.x == 0) out[0] = (float)s;
}

__global__ void sqrt_kernel(float* x) { x[0] = sqrtf(x[0]); }

template <typename T>
__global__ void __launch_bounds__(kThreads)
clip_scale_kernel(T* __restrict__ g, long long n, const float* __restrict__ total_norm, float max_norm) {
  constexpr int VEC = is16<T>::value ? 8 : 4;
  const float c = clip_coef(total_norm, max_norm);
  const long long nvec = n / VEC;
  const long long stride = (long long)gridDim.x * kThreads;
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < nvec; i += stride) {
    float v[VEC];
    load_vec<T, VEC>(g + i * VEC, v);
#pragma unroll
    for (int k = 0; k < VEC; ++k) v[k] *= c;
    store_vec<T, VEC>(g + i * VEC, v);
  }
  if (blockIdx.x == 0 && threadIdx.x < (int)(n - nvec * VEC)) {
    const long long i = nvec * VEC + threadIdx.x;
    g[i] = from_f32<T>(to_f32(g[i]) * c);
  }
}

int elems_per_cta() {
  static const int v = [] {
    const char* e = getenv("DGS_NORM_ELEMS");
    int m = e ? atoi(e) : 0;
    return m > 0 ? m : 32768;  // measured: 32 Ki elements per CTA beats 16 Ki by 4-7% (sweep_unroll.txt)
  }();
  return v;
}

int grid_for(int64_t n) {
  // one CTA per 32 Ki elements, but never fewer CTAs than fill the GPU when the input allows 4 Ki each
  // (a 25 MiB bucket would otherwise run on 200 CTAs); capped by the size of the partials scratch
  const long long per = elems_per_cta();
  long long want = (n + per - 1) / per;
  const long long fill = (long long)sm_count() * 8;
  if (want < fill) {
    const long long fine = (n + 4095) / 4096;
    want = fine < fill ? fine : fill;
  }
  if (want < 1) want = 1;
  return (int)(want < DGS_NORM_MAX_PARTIALS ? want : DGS_NORM_MAX_PARTIALS);
}

}  // namespace
}  // namespace dgs

extern "C" int dgs_l2norm_grid(int64_t n) { return n <= 0 ? 1 : dgs::grid_for(n); }

extern "C" int dgs_l2norm_partial(const void* x, int dtype, int64_t n, float* partials,
                                  dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(n >= 0, "dgs_l2norm_partial: n < 0");
  DGS_CHECK_ARG(partials != nullptr, "dgs_l2norm_partial: partials is NULL");
  DGS_CHECK_ARG(n == 0 || x != nullptr, "dgs_l2norm_partial: x is NULL");
  DGS_CHECK_ARG(aligned16(x), "dgs_l2norm_partial: x is not 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = dgs_l2norm_grid(n);
  switch (dtype) {
    case DGS_F32: sumsq_kernel<float><<<grid, kThreads, 0, st>>>((const float*)x, n, partials); break;
    case DGS_BF16: sumsq_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>((const __nv_bfloat16*)x, n, partials); break;
    case DGS_F16: sumsq_kernel<__half><<<grid, kThreads, 0, st>>>((const __half*)x, n, partials); break;
    default: return fail(DGS_EINVAL, "dgs_l2norm_partial: bad dtype %d", dtype);
  }
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_l2norm_final(const float* partials, int n_partials, float* out_norm,
                                dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(partials != nullptr && out_norm != nullptr, "dgs_l2norm_final: NULL buffer");
  DGS_CHECK_ARG(n_partials >= 0, "dgs_l2norm_final: n_partials < 0");
  norm_final_kernel<<<1, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(partials, n_partials, out_norm);
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_l2norm_final_sumsq(const float* partials, int n_partials, float* out_sumsq, dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(partials != nullptr && out_sumsq != nullptr, "dgs_l2norm_final_sumsq: NULL buffer");
  DGS_CHECK_ARG(n_partials >= 0, "dgs_l2norm_final_sumsq: n_partials < 0");
  norm_final_sumsq_kernel<<<1, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(partials, n_partials, out_sumsq);
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_sqrt_inplace(float* x, dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(x != nullptr, "dgs_sqrt_inplace: NULL buffer");
  sqrt_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(x);
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_l2norm(const void* x, int dtype, int64_t n, float* partials, float* out_norm,
                          dgs_stream_t stream) {
  int rc = dgs_l2norm_partial(x, dtype, n, partials, stream);
  if (rc != DGS_OK) return rc;
  return dgs_l2norm_final(partials, dgs_l2norm_grid(n), out_norm, stream);
}

extern "C" int dgs_clip_scale(void* g, int dtype, int64_t n, const float* total_norm, float max_norm,
                              dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(n >= 0, "dgs_clip_scale: n < 0");
  if (n == 0) return DGS_OK;
  DGS_CHECK_ARG(g != nullptr && total_norm != nullptr, "dgs_clip_scale: NULL buffer");
  DGS_CHECK_ARG(aligned16(g), "dgs_clip_scale: g is not 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  long long want = (n / 4 + kThreads - 1) / kThreads;
  if (want < 1) want = 1;
  const long long cap = (long long)sm_count() * grid_mult();
  const int grid = (int)(want < cap ? want : cap);
  switch (dtype) {
    case DGS_F32: clip_scale_kernel<float><<<grid, kThreads, 0, st>>>((float*)g, n, total_norm, max_norm); break;
    case DGS_BF16: clip_scale_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>((__nv_bfloat16*)g, n, total_norm, max_norm); break;
    case DGS_F16: clip_scale_kernel<__half><<<grid, kThreads, 0, st>>>((__half*)g, n, total_norm, max_norm); break;
    default: return fail(DGS_EINVAL, "dgs_clip_scale: bad dtype %d", dtype);
  }
  DGS_LAUNCHED();
  return DGS_OK;
}
