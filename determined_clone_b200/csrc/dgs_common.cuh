// Shared device/host helpers for libdgs.so (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "../../include/dgs.h"

namespace dgs {

// ---- host-side error plumbing (dgs_api.cu) -----------------------------------------------------
int fail(int code, const char* fmt, ...);
extern std::atomic<int64_t> g_launches;
int sm_count();  // SMs of the current device (148 on B200), cached per device
// Upper bound on the grid of a streaming kernel, in CTAs per SM (env DGS_GRID_MULT overrides, for tuning).
int grid_mult();
// Optional CUDA-event bracket around a kernel launch (dgs_timer_enable); no-ops when timing is off.
void timer_begin(cudaStream_t st);
void timer_end(cudaStream_t st);

#define DGS_CHECK_ARG(cond, ...)                     \
  do {                                               \
    if (!(cond)) return ::dgs::fail(DGS_EINVAL, __VA_ARGS__); \
  } while (0)

#define DGS_CUDA(expr)                                                                       \
  do {                                                                                       \
    cudaError_t e__ = (expr);                                                                \
    if (e__ != cudaSuccess)                                                                  \
      return ::dgs::fail(DGS_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), \
                         __FILE__, __LINE__);                                                \
  } while (0)

// call after every <<<>>>: catches launch-configuration errors and counts the launch
#define DGS_LAUNCHED()                       \
  do {                                       \
    ::dgs::g_launches.fetch_add(1);          \
    DGS_CUDA(cudaGetLastError());            \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline int dtype_size(int dt) {
  switch (dt) {
    case DGS_F32: return 4;
    case DGS_BF16: return 2;
    case DGS_F16: return 2;
    case DGS_F64: return 8;
    case DGS_I64: return 8;
    default: return 0;
  }
}

// ---- device-side vector IO -----------------------------------------------------------------
// VEC elements per thread per access.  fp32 always moves 4 elements as one 128-bit access; 16-bit
// types move 8 elements as one 128-bit access (VEC == 8) or 4 elements as one 64-bit access when they
// share a loop with an fp32 array (VEC == 4), so that every warp-level access is a contiguous run.
template <typename T>
struct is16 {
  static constexpr bool value = false;
};
template <>
struct is16<__nv_bfloat16> {
  static constexpr bool value = true;
};
template <>
struct is16<__half> {
  static constexpr bool value = true;
};

__device__ __forceinline__ float to_f32(float x) { return x; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
__device__ __forceinline__ float to_f32(__half x) { return __half2float(x); }
template <typename T>
__device__ __forceinline__ T from_f32(float x);
template <>
__device__ __forceinline__ float from_f32<float>(float x) { return x; }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float x) {
  return __float2bfloat16_rn(x);
}
template <>
__device__ __forceinline__ __half from_f32<__half>(float x) { return __float2half_rn(x); }

__device__ __forceinline__ void unpack2(uint32_t w, __nv_bfloat16*, float& lo, float& hi) {
  lo = __uint_as_float(w << 16);
  hi = __uint_as_float(w & 0xffff0000u);
}
__device__ __forceinline__ void unpack2(uint32_t w, __half*, float& lo, float& hi) {
  __half2 h = *reinterpret_cast<__half2*>(&w);
  float2 f = __half22float2(h);
  lo = f.x;
  hi = f.y;
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi, __nv_bfloat16*) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi, __half*) {
  __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

enum class Cache { Default, Stream };

template <typename T, int VEC, Cache C = Cache::Default>
__device__ __forceinline__ void load_vec(const T* __restrict__ p, float (&out)[VEC]) {
  if constexpr (!is16<T>::value) {
    static_assert(VEC == 4, "fp32 moves 4 elements per access");
    float4 v = (C == Cache::Stream) ? __ldcs(reinterpret_cast<const float4*>(p))
                                    : *reinterpret_cast<const float4*>(p);
    out[0] = v.x; out[1] = v.y; out[2] = v.z; out[3] = v.w;
  } else if constexpr (VEC == 8) {
    uint4 v = (C == Cache::Stream) ? __ldcs(reinterpret_cast<const uint4*>(p))
                                   : *reinterpret_cast<const uint4*>(p);
    unpack2(v.x, (T*)nullptr, out[0], out[1]);
    unpack2(v.y, (T*)nullptr, out[2], out[3]);
    unpack2(v.z, (T*)nullptr, out[4], out[5]);
    unpack2(v.w, (T*)nullptr, out[6], out[7]);
  } else {
    static_assert(VEC == 4, "16-bit types move 4 or 8 elements per access");
    uint2 v = (C == Cache::Stream) ? __ldcs(reinterpret_cast<const uint2*>(p))
                                   : *reinterpret_cast<const uint2*>(p);
    unpack2(v.x, (T*)nullptr, out[0], out[1]);
    unpack2(v.y, (T*)nullptr, out[2], out[3]);
  }
}

template <typename T, int VEC>
__device__ __forceinline__ void store_vec(T* __restrict__ p, const float (&in)[VEC]) {
  if constexpr (!is16<T>::value) {
    static_assert(VEC == 4, "fp32 moves 4 elements per access");
    *reinterpret_cast<float4*>(p) = make_float4(in[0], in[1], in[2], in[3]);
  } else if constexpr (VEC == 8) {
    uint4 v;
    v.x = pack2(in[0], in[1], (T*)nullptr);
    v.y = pack2(in[2], in[3], (T*)nullptr);
    v.z = pack2(in[4], in[5], (T*)nullptr);
    v.w = pack2(in[6], in[7], (T*)nullptr);
    *reinterpret_cast<uint4*>(p) = v;
  } else {
    uint2 v;
    v.x = pack2(in[0], in[1], (T*)nullptr);
    v.y = pack2(in[2], in[3], (T*)nullptr);
    *reinterpret_cast<uint2*>(p) = v;
  }
}

// ---- reductions -----------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum over a block of up to 1024 threads; result valid in thread 0.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* smem /* >= 32 entries */) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) smem[wid] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? smem[threadIdx.x] : T(0);
  if (wid == 0) v = warp_sum(v);
  return v;
}

// torch/nn/utils/clip_grad.py: clip_coef = max_norm / (total_norm + 1e-6), clamped to <= 1
__device__ __forceinline__ float clip_coef(const float* __restrict__ total_norm, float max_norm) {
  if (total_norm == nullptr) return 1.0f;
  float c = max_norm / (*total_norm + 1e-6f);
  return (c > 1.0f) ? 1.0f : c;  // NaN stays NaN, as torch.clamp(max=1.0) does
}

// Gradients that still carry a GradScaler's loss scale S: inv_scale[0] = 1/S.  The factor every gradient element
// is multiplied by is (1/S) * clip_coef(norm of the UNSCALED gradients), where total_norm[0] was measured on the
// scaled buffer.  inv_scale == nullptr multiplies by exactly 1.0f everywhere, i.e. gives clip_coef() bit for bit.
__device__ __forceinline__ float grad_coef(const float* __restrict__ total_norm, float max_norm,
                                           const float* __restrict__ inv_scale) {
  const float inv = (inv_scale != nullptr) ? *inv_scale : 1.0f;
  if (total_norm == nullptr) return inv;
  float c = max_norm / (*total_norm * inv + 1e-6f);
  c = (c > 1.0f) ? 1.0f : c;
  return c * inv;
}

}  // namespace dgs
