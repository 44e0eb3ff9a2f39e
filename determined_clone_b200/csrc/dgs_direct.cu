// Single-launch optimizer step for one GPU: flatten + scale + (global L2 norm) + clip + optimizer step in ONE kernel
// that reads the gradient tensors where autograd left them.
//
// Takes over, for a trial on one slot, everything between loss.backward() and the end of step_optimizer() at
// reference harness/determined/pytorch/_pytorch_context.py:818-924 (_average_gradients, clip_grads == clip_grad_norm_,
// optimizer.step(), optimizer.zero_grad()) -- the same lines as dgs_flatten_accumulate + dgs_l2norm + dgs_step_*, but:
//   * the averaged gradient is never written to (and re-read from) a wire buffer: the step reads the gradient tensors
//     through the segment table.  fp32 AdamW + clip: 4 (norm pass) + 28 (step) = 32 B/element instead of 40; without
//     clipping 28 (AdamW) / 20 (SGD momentum) instead of 36 / 28.
//   * one launch instead of four, which is what bounds small models (a 1-16 MB parameter set is launch-bound: the
//     whole working set sits in the 126 MB L2 and each extra launch costs more than the data movement).
// The norm needs every gradient before any parameter moves, so the kernel is launched cooperatively and crosses one
// grid-wide barrier between the sum-of-squares pass and the step pass; when every tile has its own CTA the gradient
// values stay in registers across the barrier and are read exactly once.
//
// Results match the three-kernel path: the gradient value is rounded to the parameter dtype where that path stores it
// to the wire buffer, the norm is summed from per-CTA fp32 partials in double, the update is dgs_opt.cuh's.
#include <cooperative_groups.h>

#include "dgs_common.cuh"
#include "dgs_opt.cuh"

namespace cg = cooperative_groups;

namespace dgs {
namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 2;

enum Kind { kAdam = 0, kSgdMomentum = 1, kSgdPlain = 2 };

struct DSeg {
  const void* src;
  long long dst_off;
  long long numel;
  long long tile_start;
  int vec_ok;  // src is 16-byte aligned
  int range;   // hyper-parameter row of this tensor
};

template <typename Hp>
struct DirectParams {
  DSeg seg[DGS_DIRECT_MAX_SEGS];
  Hp hp[DGS_MAX_RANGES];
  int n_segs;
  long long tiles;
};

template <int KIND>
struct HpOf {
  using type = SgdHp;
};
template <>
struct HpOf<kAdam> {
  using type = AdamHp;
};

template <typename P, typename G, bool READ_ACC>
struct DirectCfg {
  static constexpr int VEC = (is16<P>::value && is16<G>::value && !READ_ACC) ? 8 : 4;
  static constexpr int TILE = kThreads * VEC * kUnroll;
};

__device__ __forceinline__ int find_seg(const DSeg* seg, int n, long long t) {
  int lo = 0, hi = n - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (seg[mid].tile_start <= t) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// the averaged gradient of one tile, as the wire buffer would hold it: cast((acc + g) * scale) read back as fp32
template <typename P, typename G, bool READ_ACC, int VEC>
__device__ __forceinline__ void load_grad_tile(const DSeg& s, long long base, const float* __restrict__ acc, float scale,
                                               float (&v)[kUnroll][VEC]) {
  const G* __restrict__ src = static_cast<const G*>(s.src);
#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    const long long idx = base + (long long)(u * kThreads + threadIdx.x) * VEC;
    float a[VEC];
    if (idx + VEC <= s.numel && s.vec_ok) {
      load_vec<G, VEC, Cache::Stream>(src + idx, v[u]);
      if constexpr (READ_ACC) load_vec<float, 4>(acc + s.dst_off + idx, a);
    } else {
#pragma unroll
      for (int k = 0; k < VEC; ++k) {
        const bool in = idx + k < s.numel;
        v[u][k] = in ? to_f32(src[idx + k]) : 0.0f;
        if constexpr (READ_ACC) a[k] = in ? acc[s.dst_off + idx + k] : 0.0f;
      }
    }
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      float x = v[u][k];
      if constexpr (READ_ACC) x += a[k];
      v[u][k] = to_f32(from_f32<P>(x * scale));
    }
  }
}

template <typename P, typename G, int KIND, bool READ_ACC>
__global__ void __launch_bounds__(kThreads)
direct_kernel(const __grid_constant__ DirectParams<typename HpOf<KIND>::type> D, const float* __restrict__ acc, float scale,
              P* __restrict__ param, P* __restrict__ s1, P* __restrict__ s2, int clip, float max_norm,
              float* __restrict__ partials, float* __restrict__ out_norm) {
  using Cfg = DirectCfg<P, G, READ_ACC>;
  constexpr int VEC = Cfg::VEC;
  constexpr int TILE = Cfg::TILE;
  __shared__ float red_f[32];
  __shared__ double red_d[32];
  __shared__ float coef_s;

  const bool keep = D.tiles <= (long long)gridDim.x;  // one tile per CTA: its gradient stays in registers across the barrier
  float g[kUnroll][VEC];
  float coef = 1.0f;
  if (clip) {
    float ss = 0.0f;
    for (long long t = blockIdx.x; t < D.tiles; t += gridDim.x) {
      const DSeg& s = D.seg[find_seg(D.seg, D.n_segs, t)];
      load_grad_tile<P, G, READ_ACC, VEC>(s, (t - s.tile_start) * TILE, acc, scale, g);
#pragma unroll
      for (int u = 0; u < kUnroll; ++u)
#pragma unroll
        for (int k = 0; k < VEC; ++k) ss = fmaf(g[u][k], g[u][k], ss);
    }
    const float tot = block_sum(ss, red_f);
    if (threadIdx.x == 0) partials[blockIdx.x] = tot;
    __threadfence();
    cg::this_grid().sync();
    // every CTA adds the same partials in the same order: one norm, bit-identical everywhere
    double sum = 0.0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += kThreads) sum += (double)__ldcg(partials + i);
    sum = block_sum(sum, red_d);
    if (threadIdx.x == 0) {
      const float norm = (float)sqrt(sum);
      if (blockIdx.x == 0) out_norm[0] = norm;
      coef_s = clip_coef(&norm, max_norm);
    }
    __syncthreads();
    coef = coef_s;
  }

  for (long long t = blockIdx.x; t < D.tiles; t += gridDim.x) {
    const DSeg& s = D.seg[find_seg(D.seg, D.n_segs, t)];
    const long long base = (t - s.tile_start) * TILE;
    if (!(clip && keep)) load_grad_tile<P, G, READ_ACC, VEC>(s, base, acc, scale, g);
    const auto h = D.hp[s.range];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long idx = base + (long long)(u * kThreads + threadIdx.x) * VEC;
      const long long e = s.dst_off + idx;
      if (idx + VEC <= s.numel) {
        float p[VEC], a[VEC], b[VEC];
        load_vec<P, VEC>(param + e, p);
        if constexpr (KIND != kSgdPlain) load_vec<P, VEC>(s1 + e, a);
        if constexpr (KIND == kAdam) load_vec<P, VEC>(s2 + e, b);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
          if constexpr (KIND == kAdam) adam_update(p[k], a[k], b[k], g[u][k], h, coef);
          else if constexpr (KIND == kSgdMomentum) sgd_update(p[k], a[k], g[u][k], h, coef, true);
          else { float none = 0.0f; sgd_update(p[k], none, g[u][k], h, coef, false); }
        }
        store_vec<P, VEC>(param + e, p);
        if constexpr (KIND != kSgdPlain) store_vec<P, VEC>(s1 + e, a);
        if constexpr (KIND == kAdam) store_vec<P, VEC>(s2 + e, b);
      } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
          if (idx + k < s.numel) {
            float p = to_f32(param[e + k]), a = 0.0f, b = 0.0f;
            if constexpr (KIND != kSgdPlain) a = to_f32(s1[e + k]);
            if constexpr (KIND == kAdam) b = to_f32(s2[e + k]);
            if constexpr (KIND == kAdam) adam_update(p, a, b, g[u][k], h, coef);
            else sgd_update(p, a, g[u][k], h, coef, KIND == kSgdMomentum);
            param[e + k] = from_f32<P>(p);
            if constexpr (KIND != kSgdPlain) s1[e + k] = from_f32<P>(a);
            if constexpr (KIND == kAdam) s2[e + k] = from_f32<P>(b);
          }
        }
      }
    }
  }
}

template <typename P, typename G, int KIND, bool READ_ACC>
int launch_direct(DirectParams<typename HpOf<KIND>::type>& D, const dgs_seg* segs, const int32_t* seg_range, int n_segs,
                  const float* acc, float scale, void* param, void* s1, void* s2, int clip, float max_norm, float* partials,
                  int partials_capacity, float* out_norm, cudaStream_t st) {
  constexpr int TILE = DirectCfg<P, G, READ_ACC>::TILE;
  long long tiles = 0;
  int m = 0;
  for (int i = 0; i < n_segs; ++i) {
    if (segs[i].numel == 0) continue;
    DSeg& d = D.seg[m++];
    d.src = segs[i].src;
    d.dst_off = segs[i].dst_off;
    d.numel = segs[i].numel;
    d.tile_start = tiles;
    d.vec_ok = aligned16(segs[i].src) ? 1 : 0;
    d.range = seg_range[i];
    tiles += (segs[i].numel + TILE - 1) / TILE;
  }
  if (m == 0) return DGS_OK;
  D.n_segs = m;
  D.tiles = tiles;
  auto kernel = direct_kernel<P, G, KIND, READ_ACC>;
  static int per_sm = 0;  // per instantiation
  if (per_sm == 0) {
    int occ = 0;
    DGS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kThreads, 0));
    if (occ < 1) return fail(DGS_ECUDA, "dgs_direct_step: the kernel does not fit on an SM");
    per_sm = occ;
  }
  P* pp = static_cast<P*>(param);
  P* p1 = static_cast<P*>(s1);
  P* p2 = static_cast<P*>(s2);
  if (!clip) {
    // no norm, no grid barrier: an ordinary launch, one tile per CTA (measured faster than a persistent grid for
    // streaming read/write kernels on B200, see grid_mult() in dgs_api.cu)
    const long long cap = (long long)sm_count() * grid_mult();
    const int grid = (int)(tiles < cap ? tiles : cap);
    timer_begin(st);
    kernel<<<grid, kThreads, 0, st>>>(D, acc, scale, pp, p1, p2, 0, max_norm, partials, out_norm);
    timer_end(st);
    DGS_LAUNCHED();
    return DGS_OK;
  }
  long long grid = (long long)sm_count() * per_sm;  // a cooperative launch needs every CTA resident
  if (grid > tiles) grid = tiles;
  if (grid > partials_capacity) grid = partials_capacity;
  void* args[] = {&D, &acc, &scale, &pp, &p1, &p2, &clip, &max_norm, &partials, &out_norm};
  timer_begin(st);
  DGS_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kernel), dim3((unsigned)grid), dim3(kThreads), args, 0, st));
  timer_end(st);
  DGS_LAUNCHED();
  return DGS_OK;
}

int check_direct(const char* who, const dgs_seg* segs, const int32_t* seg_range, int n_segs, int n_ranges, const void* acc,
                 void* param, int clip, const float* partials, int partials_capacity, const float* out_norm) {
  DGS_CHECK_ARG(n_segs >= 0 && n_segs <= DGS_DIRECT_MAX_SEGS, "%s: n_segs %d outside [0, %d]", who, n_segs, DGS_DIRECT_MAX_SEGS);
  DGS_CHECK_ARG(n_ranges >= 1 && n_ranges <= DGS_MAX_RANGES, "%s: n_ranges %d outside [1, %d]", who, n_ranges, DGS_MAX_RANGES);
  DGS_CHECK_ARG(n_segs == 0 || (segs != nullptr && seg_range != nullptr), "%s: segs / seg_range is NULL", who);
  DGS_CHECK_ARG(param != nullptr && aligned16(param) && aligned16(acc), "%s: param is NULL or a buffer is misaligned", who);
  DGS_CHECK_ARG(!clip || (partials != nullptr && partials_capacity >= 1 && out_norm != nullptr),
                "%s: clipping needs partials scratch and out_norm", who);
  for (int i = 0; i < n_segs; ++i) {
    DGS_CHECK_ARG(segs[i].numel >= 0 && segs[i].dst_off >= 0 && (segs[i].dst_off & 7) == 0,
                  "%s: seg %d: negative size or dst_off not a multiple of 8", who, i);
    DGS_CHECK_ARG(segs[i].numel == 0 || segs[i].src != nullptr, "%s: seg %d src is NULL", who, i);
    DGS_CHECK_ARG(seg_range[i] >= 0 && seg_range[i] < n_ranges, "%s: seg %d names range %d of %d", who, i, seg_range[i], n_ranges);
  }
  return DGS_OK;
}

}  // namespace
}  // namespace dgs

#define DGS_DIRECT_DISPATCH(KIND, S1, S2)                                                                                     \
  if (state_dtype == DGS_F32 && grad_dtype == DGS_F32) {                                                                      \
    if (acc) return launch_direct<float, float, KIND, true>(D, segs, seg_range, n_segs, acc, scale, param, S1, S2, clip,      \
                                                            max_norm, partials, partials_capacity, out_norm, st);             \
    return launch_direct<float, float, KIND, false>(D, segs, seg_range, n_segs, acc, scale, param, S1, S2, clip, max_norm,    \
                                                    partials, partials_capacity, out_norm, st);                               \
  }                                                                                                                           \
  if (state_dtype == DGS_BF16 && grad_dtype == DGS_BF16) {                                                                    \
    if (acc) return launch_direct<__nv_bfloat16, __nv_bfloat16, KIND, true>(D, segs, seg_range, n_segs, acc, scale, param,    \
                                                                            S1, S2, clip, max_norm, partials,                 \
                                                                            partials_capacity, out_norm, st);                 \
    return launch_direct<__nv_bfloat16, __nv_bfloat16, KIND, false>(D, segs, seg_range, n_segs, acc, scale, param, S1, S2,    \
                                                                    clip, max_norm, partials, partials_capacity, out_norm,    \
                                                                    st);                                                      \
  }                                                                                                                           \
  if (state_dtype == DGS_F16 && grad_dtype == DGS_F16) {                                                                      \
    if (acc) return launch_direct<__half, __half, KIND, true>(D, segs, seg_range, n_segs, acc, scale, param, S1, S2, clip,    \
                                                              max_norm, partials, partials_capacity, out_norm, st);           \
    return launch_direct<__half, __half, KIND, false>(D, segs, seg_range, n_segs, acc, scale, param, S1, S2, clip, max_norm,  \
                                                      partials, partials_capacity, out_norm, st);                             \
  }

extern "C" int dgs_direct_step_adam(const dgs_seg* segs, const int32_t* seg_range, int n_segs, int grad_dtype,
                                    const float* acc, float scale, void* param, void* exp_avg, void* exp_avg_sq,
                                    int state_dtype, const dgs_adam_range* ranges, int n_ranges, int clip, float max_norm,
                                    float* partials, int partials_capacity, float* out_norm, dgs_stream_t stream) {
  using namespace dgs;
  if (int rc = check_direct("dgs_direct_step_adam", segs, seg_range, n_segs, n_ranges, acc, param, clip, partials,
                            partials_capacity, out_norm))
    return rc;
  DGS_CHECK_ARG(ranges != nullptr && exp_avg && exp_avg_sq && aligned16(exp_avg) && aligned16(exp_avg_sq),
                "dgs_direct_step_adam: NULL or misaligned state");
  static thread_local DirectParams<AdamHp> D;  // ~22 KB: too large for a comfortable stack frame
  for (int i = 0; i < n_ranges; ++i) fill_adam_hp(ranges[i], D.hp[i]);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DGS_DIRECT_DISPATCH(kAdam, exp_avg, exp_avg_sq)
  return fail(DGS_EUNSUPPORTED, "dgs_direct_step_adam: state dtype %d with grad dtype %d", state_dtype, grad_dtype);
}

extern "C" int dgs_direct_step_sgd(const dgs_seg* segs, const int32_t* seg_range, int n_segs, int grad_dtype,
                                   const float* acc, float scale, void* param, void* momentum_buf, int state_dtype,
                                   const dgs_sgd_range* ranges, int n_ranges, int clip, float max_norm, float* partials,
                                   int partials_capacity, float* out_norm, dgs_stream_t stream) {
  using namespace dgs;
  if (int rc = check_direct("dgs_direct_step_sgd", segs, seg_range, n_segs, n_ranges, acc, param, clip, partials,
                            partials_capacity, out_norm))
    return rc;
  DGS_CHECK_ARG(ranges != nullptr && aligned16(momentum_buf), "dgs_direct_step_sgd: NULL ranges or misaligned momentum_buf");
  bool any_mom = false, all_mom = true;
  for (int i = 0; i < n_ranges; ++i) {
    if (ranges[i].momentum != 0.0) any_mom = true; else all_mom = false;
  }
  DGS_CHECK_ARG(!any_mom || (all_mom && momentum_buf != nullptr),
                "dgs_direct_step_sgd: ranges with and without momentum must go in separate calls; momentum needs momentum_buf");
  static thread_local DirectParams<SgdHp> D;
  for (int i = 0; i < n_ranges; ++i) fill_sgd_hp(ranges[i], D.hp[i]);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void* none = nullptr;
  if (any_mom) {
    DGS_DIRECT_DISPATCH(kSgdMomentum, momentum_buf, none)
  } else {
    DGS_DIRECT_DISPATCH(kSgdPlain, none, none)
  }
  return fail(DGS_EUNSUPPORTED, "dgs_direct_step_sgd: state dtype %d with grad dtype %d", state_dtype, grad_dtype);
}
