// Fused global-norm clip + optimizer step over flat param / state / grad buffers.
//
// Takes over clip_grads(parameters) + optimizer.step() + optimizer.zero_grad() of the reference at
// harness/determined/pytorch/_pytorch_context.py:902-924 for torch.optim.Adam/AdamW and
// torch.optim.SGD.  Arithmetic follows torch 2.11 torch/optim/adam.py (_multi_tensor_adam) and
// torch/optim/sgd.py (_multi_tensor_sgd) op by op, in fp32, with one rounding at the store for
// 16-bit params/state.
//
// HBM-bound: AdamW moves 28 B/element in fp32 (read g,p,m,v; write p,m,v), SGD-momentum 20 B.  Each
// thread keeps 3-4 independent 16 B loads in flight (one per array); every warp access is contiguous; one tile per
// CTA (measured faster on B200 than a persistent grid-stride loop, see grid_mult() in dgs_api.cu).
#include "dgs_common.cuh"
#include "dgs_opt.cuh"

#include <cstdlib>

namespace dgs {
namespace {

constexpr int kThreads = 256;

template <typename Hp>
struct Table {
  Hp r[DGS_MAX_RANGES];
  int n;
};

// Destinations of the updated parameters.  Plain step: the local buffer only.  Push step (dgs_step_*_push, the
// all-gather half of the peer exchange, dgs_peer.cu): the same element offset in every rank's parameter buffer,
// peer-mapped over NVLink; dst[own rank] is the local buffer the kernel also reads.
struct ParamDst {
  void* dst[DGS_PEER_MAX];
  int n;
};

template <bool PUSH, typename P, int VEC>
__device__ __forceinline__ void store_param(P* __restrict__ local, const ParamDst& D, long long i, const float (&v)[VEC]) {
  if constexpr (PUSH) {
    // destinations in an order rotated per CTA: the CTAs in flight then do not all address the same link at once
    const int r0 = (int)(blockIdx.x % (unsigned)D.n);
#pragma unroll
    for (int k = 0; k < DGS_PEER_MAX; ++k)
      if (k < D.n) {
        const int r = (r0 + k < D.n) ? r0 + k : r0 + k - D.n;
        store_vec<P, VEC>(static_cast<P*>(D.dst[r]) + i, v);
      }
  } else {
    store_vec<P, VEC>(local + i, v);
  }
}
template <bool PUSH, typename P>
__device__ __forceinline__ void store_param1(P* __restrict__ local, const ParamDst& D, long long e, float v) {
  if constexpr (PUSH) {
    for (int r = 0; r < D.n; ++r) static_cast<P*>(D.dst[r])[e] = from_f32<P>(v);
  } else {
    local[e] = from_f32<P>(v);
  }
}

template <typename Hp>
__device__ __forceinline__ int find_range(const Table<Hp>& T, long long e) {
  for (int i = 0; i < T.n; ++i)
    if (e >= T.r[i].begin && e < T.r[i].end) return i;
  return -1;
}

template <typename P, typename G, int kUnroll>
struct StepCfg {
  static constexpr int VEC = (is16<P>::value && is16<G>::value) ? 8 : 4;
  static constexpr int TILE = kThreads * VEC * kUnroll;
};

inline int step_unroll() {
  static const int v = [] {
    const char* e = getenv("DGS_STEP_UNROLL");
    return e ? atoi(e) : 0;
  }();
  return v;
}

template <typename P, typename G, int kUnroll, bool PUSH>
__global__ void __launch_bounds__(kThreads)
adam_kernel(P* __restrict__ param, P* __restrict__ ea, P* __restrict__ eas, const G* __restrict__ grad,
            long long n, const __grid_constant__ Table<AdamHp> T,
            const float* __restrict__ total_norm, float max_norm, const float* __restrict__ inv_scale,
            const __grid_constant__ ParamDst D) {
  constexpr int VEC = StepCfg<P, G, kUnroll>::VEC;
  constexpr int TILE = StepCfg<P, G, kUnroll>::TILE;
  const float clip = grad_coef(total_norm, max_norm, inv_scale);
  const long long tiles = (n + TILE - 1) / TILE;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long e0 = t * TILE;
    // first range that ends after the tile start (ranges are sorted and disjoint)
    int r = 0;
    while (r < T.n && T.r[r].end <= e0) ++r;
    if (r == T.n) break;
    // a tile that no range touches (another rank's slice when the ranges are this rank's slices of every bucket) is
    // skipped outright: left to the per-element path below it cost ~2 us of issue slots each, 1 ms per step on GPT-2
    // medium at 8 GPUs (profiles/r2_sweep_n8_gpt2.txt)
    if (T.r[r].begin >= e0 + TILE) continue;
    if (T.r[r].begin <= e0 && e0 + TILE <= T.r[r].end) {
      const AdamHp h = T.r[r];
      float g[kUnroll][VEC], p[kUnroll][VEC], m[kUnroll][VEC], v[kUnroll][VEC];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = e0 + (long long)(u * kThreads + threadIdx.x) * VEC;
        load_vec<G, VEC, Cache::Stream>(grad + i, g[u]);
        load_vec<P, VEC>(param + i, p[u]);
        load_vec<P, VEC>(ea + i, m[u]);
        load_vec<P, VEC>(eas + i, v[u]);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = e0 + (long long)(u * kThreads + threadIdx.x) * VEC;
#pragma unroll
        for (int k = 0; k < VEC; ++k) adam_update(p[u][k], m[u][k], v[u][k], g[u][k], h, clip);
        store_param<PUSH, P, VEC>(param, D, i, p[u]);
        store_vec<P, VEC>(ea + i, m[u]);
        store_vec<P, VEC>(eas + i, v[u]);
      }
    } else {
      // tile straddles a range boundary or the end of the buffer: scalar, per-element lookup
      const long long e1 = (e0 + TILE < n) ? e0 + TILE : n;
      for (long long e = e0 + threadIdx.x; e < e1; e += kThreads) {
        const int q = find_range(T, e);
        if (q < 0) continue;
        float p = to_f32(param[e]), m = to_f32(ea[e]), v = to_f32(eas[e]);
        adam_update(p, m, v, to_f32(grad[e]), T.r[q], clip);
        store_param1<PUSH, P>(param, D, e, p);
        ea[e] = from_f32<P>(m);
        eas[e] = from_f32<P>(v);
      }
    }
  }
}

template <typename P, typename G, bool HAS_BUF, int kUnroll, bool PUSH>
__global__ void __launch_bounds__(kThreads)
sgd_kernel(P* __restrict__ param, P* __restrict__ buf, const G* __restrict__ grad, long long n,
           const __grid_constant__ Table<SgdHp> T, const float* __restrict__ total_norm,
           float max_norm, const float* __restrict__ inv_scale, const __grid_constant__ ParamDst D) {
  constexpr int VEC = StepCfg<P, G, kUnroll>::VEC;
  constexpr int TILE = StepCfg<P, G, kUnroll>::TILE;
  const float clip = grad_coef(total_norm, max_norm, inv_scale);
  const long long tiles = (n + TILE - 1) / TILE;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const long long e0 = t * TILE;
    int r = 0;
    while (r < T.n && T.r[r].end <= e0) ++r;
    if (r == T.n) break;
    // a tile that no range touches (another rank's slice when the ranges are this rank's slices of every bucket) is
    // skipped outright: left to the per-element path below it cost ~2 us of issue slots each, 1 ms per step on GPT-2
    // medium at 8 GPUs (profiles/r2_sweep_n8_gpt2.txt)
    if (T.r[r].begin >= e0 + TILE) continue;
    if (T.r[r].begin <= e0 && e0 + TILE <= T.r[r].end) {
      const SgdHp h = T.r[r];
      const bool read_buf = HAS_BUF && !(h.flags & DGS_SGD_FIRST_STEP);
      float g[kUnroll][VEC], p[kUnroll][VEC], b[kUnroll][VEC];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = e0 + (long long)(u * kThreads + threadIdx.x) * VEC;
        load_vec<G, VEC, Cache::Stream>(grad + i, g[u]);
        load_vec<P, VEC>(param + i, p[u]);
        if (read_buf) {
          load_vec<P, VEC>(buf + i, b[u]);
        } else {
#pragma unroll
          for (int k = 0; k < VEC; ++k) b[u][k] = 0.0f;
        }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = e0 + (long long)(u * kThreads + threadIdx.x) * VEC;
#pragma unroll
        for (int k = 0; k < VEC; ++k) sgd_update(p[u][k], b[u][k], g[u][k], h, clip, HAS_BUF);
        store_param<PUSH, P, VEC>(param, D, i, p[u]);
        if constexpr (HAS_BUF) store_vec<P, VEC>(buf + i, b[u]);
      }
    } else {
      const long long e1 = (e0 + TILE < n) ? e0 + TILE : n;
      for (long long e = e0 + threadIdx.x; e < e1; e += kThreads) {
        const int q = find_range(T, e);
        if (q < 0) continue;
        float p = to_f32(param[e]);
        float b = 0.0f;
        if (HAS_BUF && !(T.r[q].flags & DGS_SGD_FIRST_STEP)) b = to_f32(buf[e]);
        sgd_update(p, b, to_f32(grad[e]), T.r[q], clip, HAS_BUF);
        store_param1<PUSH, P>(param, D, e, p);
        if constexpr (HAS_BUF) buf[e] = from_f32<P>(b);
      }
    }
  }
}

inline int tile_grid(long long tiles) {
  const long long cap = (long long)sm_count() * grid_mult();
  if (tiles < 1) tiles = 1;
  return (int)(tiles < cap ? tiles : cap);
}

template <typename Range>
int check_ranges(const char* who, const Range* ranges, int n_ranges, int64_t n) {
  DGS_CHECK_ARG(n_ranges >= 0 && n_ranges <= DGS_MAX_RANGES, "%s: n_ranges %d outside [0, %d]", who,
                n_ranges, DGS_MAX_RANGES);
  DGS_CHECK_ARG(n_ranges == 0 || ranges != nullptr, "%s: ranges is NULL", who);
  int64_t prev = 0;
  for (int i = 0; i < n_ranges; ++i) {
    DGS_CHECK_ARG(ranges[i].begin >= prev && ranges[i].end >= ranges[i].begin && ranges[i].end <= n,
                  "%s: range %d [%lld, %lld) is unsorted, overlapping or outside [0, %lld)", who, i,
                  (long long)ranges[i].begin, (long long)ranges[i].end, (long long)n);
    DGS_CHECK_ARG((ranges[i].begin & 7) == 0 && ((ranges[i].end & 7) == 0 || ranges[i].end == n),
                  "%s: range %d bounds must be multiples of 8 elements", who, i);
    prev = ranges[i].end;
  }
  return DGS_OK;
}

const ParamDst kLocalOnly = {{nullptr}, 0};

template <typename P, typename G, int U>
int launch_adam_u(void* param, void* ea, void* eas, const void* grad, int64_t n, const Table<AdamHp>& T,
                  const float* tn, float max_norm, const float* inv, const ParamDst* push, cudaStream_t st) {
  const long long tiles = (n + StepCfg<P, G, U>::TILE - 1) / StepCfg<P, G, U>::TILE;
  const int grid = tile_grid(tiles);
  timer_begin(st);
  if (push != nullptr) {
    adam_kernel<P, G, U, true><<<grid, kThreads, 0, st>>>((P*)param, (P*)ea, (P*)eas, (const G*)grad, n, T, tn, max_norm, inv, *push);
  } else {
    adam_kernel<P, G, U, false><<<grid, kThreads, 0, st>>>((P*)param, (P*)ea, (P*)eas, (const G*)grad, n, T, tn, max_norm, inv, kLocalOnly);
  }
  timer_end(st);
  DGS_LAUNCHED();
  return DGS_OK;
}

template <typename P, typename G>
int launch_adam(void* param, void* ea, void* eas, const void* grad, int64_t n, const Table<AdamHp>& T,
                const float* tn, float max_norm, const float* inv, const ParamDst* push, cudaStream_t st) {
  // one vector per array per thread (4 loads in flight) measured best; DGS_STEP_UNROLL=2|4 re-runs the sweep
  // the pushing variant: two vectors per thread measured best at 8 GPUs (profiles/r2_sweep_n8_push_variants.txt)
  const int u = step_unroll() ? step_unroll() : (push != nullptr ? 2 : 1);
  if (u == 2) return launch_adam_u<P, G, 2>(param, ea, eas, grad, n, T, tn, max_norm, inv, push, st);
  if (u == 4) return launch_adam_u<P, G, 4>(param, ea, eas, grad, n, T, tn, max_norm, inv, push, st);
  return launch_adam_u<P, G, 1>(param, ea, eas, grad, n, T, tn, max_norm, inv, push, st);
}

template <typename P, typename G, int U>
int launch_sgd_u(void* param, void* buf, const void* grad, int64_t n, const Table<SgdHp>& T,
                 const float* tn, float max_norm, const float* inv, const ParamDst* push, cudaStream_t st) {
  const long long tiles = (n + StepCfg<P, G, U>::TILE - 1) / StepCfg<P, G, U>::TILE;
  const int grid = tile_grid(tiles);
  timer_begin(st);
  if (push != nullptr) {
    if (buf != nullptr) {
      sgd_kernel<P, G, true, U, true><<<grid, kThreads, 0, st>>>((P*)param, (P*)buf, (const G*)grad, n, T, tn, max_norm, inv, *push);
    } else {
      sgd_kernel<P, G, false, U, true><<<grid, kThreads, 0, st>>>((P*)param, nullptr, (const G*)grad, n, T, tn, max_norm, inv, *push);
    }
  } else if (buf != nullptr) {
    sgd_kernel<P, G, true, U, false><<<grid, kThreads, 0, st>>>((P*)param, (P*)buf, (const G*)grad, n, T, tn, max_norm, inv, kLocalOnly);
  } else {
    sgd_kernel<P, G, false, U, false><<<grid, kThreads, 0, st>>>((P*)param, nullptr, (const G*)grad, n, T, tn, max_norm, inv, kLocalOnly);
  }
  timer_end(st);
  DGS_LAUNCHED();
  return DGS_OK;
}

template <typename P, typename G>
int launch_sgd(void* param, void* buf, const void* grad, int64_t n, const Table<SgdHp>& T,
               const float* tn, float max_norm, const float* inv, const ParamDst* push, cudaStream_t st) {
  const int u = step_unroll() ? step_unroll() : (push != nullptr ? 2 : 1);
  if (u == 2) return launch_sgd_u<P, G, 2>(param, buf, grad, n, T, tn, max_norm, inv, push, st);
  if (u == 4) return launch_sgd_u<P, G, 4>(param, buf, grad, n, T, tn, max_norm, inv, push, st);
  return launch_sgd_u<P, G, 1>(param, buf, grad, n, T, tn, max_norm, inv, push, st);
}

// the push variants' destination table: every rank's parameter pointer for the same element range, [rank] local
int fill_dst(const char* who, void* param, void* const* param_peers, int nranks, int rank, ParamDst* D) {
  DGS_CHECK_ARG(nranks >= 2 && nranks <= DGS_PEER_MAX && rank >= 0 && rank < nranks, "%s: rank %d of %d (at most %d ranks)", who,
                rank, nranks, DGS_PEER_MAX);
  DGS_CHECK_ARG(param_peers != nullptr, "%s: param_peers is NULL", who);
  for (int r = 0; r < DGS_PEER_MAX; ++r) D->dst[r] = nullptr;
  for (int r = 0; r < nranks; ++r) {
    DGS_CHECK_ARG(param_peers[r] != nullptr && aligned16(param_peers[r]), "%s: parameter pointer of rank %d is NULL or misaligned", who, r);
    D->dst[r] = param_peers[r];
  }
  DGS_CHECK_ARG(param_peers[rank] == param, "%s: param_peers[rank] must be the local parameter pointer", who);
  D->n = nranks;
  return DGS_OK;
}

int step_adam_impl(void* param, void* exp_avg, void* exp_avg_sq, int state_dtype, const void* grad, int grad_dtype,
                   int64_t n, const dgs_adam_range* ranges, int n_ranges, const float* total_norm, float max_norm,
                   const float* inv_scale, const ParamDst* push, cudaStream_t st) {
  DGS_CHECK_ARG(n >= 0, "dgs_step_adam: n < 0");
  if (int rc = check_ranges("dgs_step_adam", ranges, n_ranges, n)) return rc;
  if (n == 0 || n_ranges == 0) return DGS_OK;
  DGS_CHECK_ARG(param && exp_avg && exp_avg_sq && grad, "dgs_step_adam: NULL buffer");
  DGS_CHECK_ARG(aligned16(param) && aligned16(exp_avg) && aligned16(exp_avg_sq) && aligned16(grad),
                "dgs_step_adam: buffers must be 16-byte aligned");
  Table<AdamHp> T;
  T.n = n_ranges;
  for (int i = 0; i < n_ranges; ++i) fill_adam_hp(ranges[i], T.r[i]);
#define DGS_ADAM_CASE(PD, PT, GD, GT) \
  if (state_dtype == PD && grad_dtype == GD) \
    return launch_adam<PT, GT>(param, exp_avg, exp_avg_sq, grad, n, T, total_norm, max_norm, inv_scale, push, st);
  DGS_ADAM_CASE(DGS_F32, float, DGS_F32, float)
  DGS_ADAM_CASE(DGS_F32, float, DGS_BF16, __nv_bfloat16)
  DGS_ADAM_CASE(DGS_F32, float, DGS_F16, __half)
  DGS_ADAM_CASE(DGS_BF16, __nv_bfloat16, DGS_F32, float)
  DGS_ADAM_CASE(DGS_BF16, __nv_bfloat16, DGS_BF16, __nv_bfloat16)
  DGS_ADAM_CASE(DGS_F16, __half, DGS_F32, float)
  DGS_ADAM_CASE(DGS_F16, __half, DGS_F16, __half)
#undef DGS_ADAM_CASE
  return fail(DGS_EUNSUPPORTED, "dgs_step_adam: state dtype %d with grad dtype %d", state_dtype, grad_dtype);
}

int step_sgd_impl(void* param, void* momentum_buf, int state_dtype, const void* grad, int grad_dtype, int64_t n,
                  const dgs_sgd_range* ranges, int n_ranges, const float* total_norm, float max_norm,
                  const float* inv_scale, const ParamDst* push, cudaStream_t st) {
  DGS_CHECK_ARG(n >= 0, "dgs_step_sgd: n < 0");
  if (int rc = check_ranges("dgs_step_sgd", ranges, n_ranges, n)) return rc;
  if (n == 0 || n_ranges == 0) return DGS_OK;
  DGS_CHECK_ARG(param && grad, "dgs_step_sgd: NULL buffer");
  DGS_CHECK_ARG(aligned16(param) && aligned16(grad) && aligned16(momentum_buf),
                "dgs_step_sgd: buffers must be 16-byte aligned");
  Table<SgdHp> T;
  T.n = n_ranges;
  for (int i = 0; i < n_ranges; ++i) {
    const dgs_sgd_range& r = ranges[i];
    DGS_CHECK_ARG(r.momentum == 0.0 || momentum_buf != nullptr,
                  "dgs_step_sgd: range %d has momentum but momentum_buf is NULL", i);
    fill_sgd_hp(r, T.r[i]);
  }
  // a buffer is only touched by ranges with momentum != 0; mixed tables are split by the caller
  bool any_mom = false, all_mom = true;
  for (int i = 0; i < n_ranges; ++i) {
    if (ranges[i].momentum != 0.0) any_mom = true; else all_mom = false;
  }
  DGS_CHECK_ARG(!any_mom || all_mom, "dgs_step_sgd: ranges with and without momentum must go in separate calls");
  void* buf = any_mom ? momentum_buf : nullptr;
#define DGS_SGD_CASE(PD, PT, GD, GT) \
  if (state_dtype == PD && grad_dtype == GD) \
    return launch_sgd<PT, GT>(param, buf, grad, n, T, total_norm, max_norm, inv_scale, push, st);
  DGS_SGD_CASE(DGS_F32, float, DGS_F32, float)
  DGS_SGD_CASE(DGS_F32, float, DGS_BF16, __nv_bfloat16)
  DGS_SGD_CASE(DGS_F32, float, DGS_F16, __half)
  DGS_SGD_CASE(DGS_BF16, __nv_bfloat16, DGS_F32, float)
  DGS_SGD_CASE(DGS_BF16, __nv_bfloat16, DGS_BF16, __nv_bfloat16)
  DGS_SGD_CASE(DGS_F16, __half, DGS_F32, float)
  DGS_SGD_CASE(DGS_F16, __half, DGS_F16, __half)
#undef DGS_SGD_CASE
  return fail(DGS_EUNSUPPORTED, "dgs_step_sgd: state dtype %d with grad dtype %d", state_dtype, grad_dtype);
}

}  // namespace
}  // namespace dgs

extern "C" int dgs_step_adam(void* param, void* exp_avg, void* exp_avg_sq, int state_dtype,
                             const void* grad, int grad_dtype, int64_t n,
                             const dgs_adam_range* ranges, int n_ranges, const float* total_norm,
                             float max_norm, dgs_stream_t stream) {
  return dgs::step_adam_impl(param, exp_avg, exp_avg_sq, state_dtype, grad, grad_dtype, n, ranges, n_ranges, total_norm,
                             max_norm, nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

extern "C" int dgs_step_adam_scaled(void* param, void* exp_avg, void* exp_avg_sq, int state_dtype,
                                    const void* grad, int grad_dtype, int64_t n,
                                    const dgs_adam_range* ranges, int n_ranges, const float* total_norm,
                                    float max_norm, const float* inv_scale, dgs_stream_t stream) {
  return dgs::step_adam_impl(param, exp_avg, exp_avg_sq, state_dtype, grad, grad_dtype, n, ranges, n_ranges, total_norm,
                             max_norm, inv_scale, nullptr, static_cast<cudaStream_t>(stream));
}

extern "C" int dgs_step_adam_push(void* param, void* exp_avg, void* exp_avg_sq, int state_dtype, const void* grad,
                                  int grad_dtype, int64_t n, const dgs_adam_range* ranges, int n_ranges,
                                  const float* total_norm, float max_norm, void* const* param_peers, int nranks,
                                  int rank, dgs_stream_t stream) {
  dgs::ParamDst D;
  if (int rc = dgs::fill_dst("dgs_step_adam_push", param, param_peers, nranks, rank, &D)) return rc;
  return dgs::step_adam_impl(param, exp_avg, exp_avg_sq, state_dtype, grad, grad_dtype, n, ranges, n_ranges, total_norm,
                             max_norm, nullptr, &D, static_cast<cudaStream_t>(stream));
}

extern "C" int dgs_step_sgd(void* param, void* momentum_buf, int state_dtype, const void* grad,
                            int grad_dtype, int64_t n, const dgs_sgd_range* ranges, int n_ranges,
                            const float* total_norm, float max_norm, dgs_stream_t stream) {
  return dgs::step_sgd_impl(param, momentum_buf, state_dtype, grad, grad_dtype, n, ranges, n_ranges, total_norm,
                            max_norm, nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

extern "C" int dgs_step_sgd_scaled(void* param, void* momentum_buf, int state_dtype, const void* grad,
                                   int grad_dtype, int64_t n, const dgs_sgd_range* ranges, int n_ranges,
                                   const float* total_norm, float max_norm, const float* inv_scale,
                                   dgs_stream_t stream) {
  return dgs::step_sgd_impl(param, momentum_buf, state_dtype, grad, grad_dtype, n, ranges, n_ranges, total_norm,
                            max_norm, inv_scale, nullptr, static_cast<cudaStream_t>(stream));
}

extern "C" int dgs_step_sgd_push(void* param, void* momentum_buf, int state_dtype, const void* grad, int grad_dtype,
                                 int64_t n, const dgs_sgd_range* ranges, int n_ranges, const float* total_norm,
                                 float max_norm, void* const* param_peers, int nranks, int rank, dgs_stream_t stream) {
  dgs::ParamDst D;
  if (int rc = dgs::fill_dst("dgs_step_sgd_push", param, param_peers, nranks, rank, &D)) return rc;
  return dgs::step_sgd_impl(param, momentum_buf, state_dtype, grad, grad_dtype, n, ranges, n_ranges, total_norm,
                            max_norm, nullptr, &D, static_cast<cudaStream_t>(stream));
}
