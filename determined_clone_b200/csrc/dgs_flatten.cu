// flatten / accumulate / scale / cast: many gradient tensors -> one persistent flat bucket.
//
// Takes over autograd's `grad += new` under DDP.no_sync() (reference
// harness/determined/pytorch/_pytorch_context.py:802-812), DDP's grad/N copy into its buckets
// (:813-816) and the per-parameter div_(A) of _average_gradients (:818-825).
//
// HBM-bound, no reuse: one pass, every warp-level access a contiguous 512 B (fp32) run, 2 independent
// 16 B accesses per array per thread in flight, one tile per CTA (see grid_mult() in dgs_api.cu).
#include "dgs_common.cuh"

#include <cstdlib>

namespace dgs {
namespace {

constexpr int kThreads = 256;
constexpr int kMaxSegs = 64;

struct SegDev {
  const void* src;
  long long dst_off;
  long long numel;
  long long tile_start;  // index of this segment's first tile
  int vec_ok;            // src is 16-byte aligned
  int pad_;
};

struct FlatParams {
  SegDev seg[kMaxSegs];
  int n;
  long long tiles;
};

// Push mode (dgs_flatten_push, the reduce-scatter half of the peer exchange, dgs_peer.cu): the bucket [begin, begin +
// n*span) is owned slice by slice by the n ranks; a value for flat element f of slice j is stored into rank j's
// wire buffer (peer-mapped over NVLink) at row `rank` of that bucket: element f + (rank - j) * span.
struct PushDst {
  void* wire[DGS_PEER_MAX];
  long long begin, span;
  long long per;  // ceil(tiles / n): tiles are visited owner-interleaved, see flatten_kernel
  int n, rank;
};

template <typename W>
__device__ __forceinline__ W* push_ptr(const PushDst& D, long long f) {
  const long long off = f - D.begin;
  int j = 0;
#pragma unroll
  for (int r = 1; r < DGS_PEER_MAX; ++r) j += (r < D.n && off >= r * D.span) ? 1 : 0;
  return static_cast<W*>(D.wire[j]) + f + (long long)(D.rank - j) * D.span;
}

template <typename G, typename W, bool READ_ACC, bool WRITE_WIRE, int kUnroll>
struct FlatCfg {
  // 8 elements per access only when every array in the loop is 16-bit
  static constexpr int VEC = (is16<G>::value && WRITE_WIRE && !READ_ACC && is16<W>::value) ? 8 : 4;
  static constexpr int TILE = kThreads * VEC * kUnroll;
};

template <typename G, typename W, bool READ_ACC, bool WRITE_WIRE, int kUnroll, bool PUSH = false>
__global__ void __launch_bounds__(kThreads)
flatten_kernel(const __grid_constant__ FlatParams P, float* __restrict__ acc, W* __restrict__ wire,
               float scale, const __grid_constant__ PushDst D) {
  using Cfg = FlatCfg<G, W, READ_ACC, WRITE_WIRE, kUnroll>;
  constexpr int VEC = Cfg::VEC;
  constexpr int TILE = Cfg::TILE;

  // Push mode visits the tiles owner-interleaved and rotated by rank: tiles in address order would make every rank
  // store to owner 0 first, then owner 1, ... -- N senders into one GPU's NVLink ingress while the other links idle
  // (measured at 8 GPUs: 3x slower than NCCL).  CTAs that run at the same time now target all owners evenly, and rank
  // r starts with owner r+1.
  const long long n_visit = PUSH ? D.per * D.n : P.tiles;
  for (long long i = blockIdx.x; i < n_visit; i += gridDim.x) {
    long long t = i;
    if constexpr (PUSH) {
      const long long q = i / D.n;
      const int o = (int)((i % D.n + D.rank + 1) % D.n);
      t = o * D.per + q;
      if (t >= P.tiles) continue;
    }
    // segment lookup: last segment whose tile_start <= t (block-uniform)
    int lo = 0, hi = P.n - 1;
    while (lo < hi) {
      int mid = (lo + hi + 1) >> 1;
      if (P.seg[mid].tile_start <= t) lo = mid; else hi = mid - 1;
    }
    const G* __restrict__ src = static_cast<const G*>(P.seg[lo].src);
    const long long numel = P.seg[lo].numel;
    const long long dst0 = P.seg[lo].dst_off;
    const long long base = (t - P.seg[lo].tile_start) * TILE;
    const bool vec_ok = P.seg[lo].vec_ok != 0;

    float g[kUnroll][VEC];
    float a[kUnroll][VEC];
    long long idx[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      idx[u] = base + (long long)(u * kThreads + threadIdx.x) * VEC;
      if (idx[u] + VEC <= numel && vec_ok) {
        load_vec<G, VEC, Cache::Stream>(src + idx[u], g[u]);
        if constexpr (READ_ACC) load_vec<float, 4>(acc + dst0 + idx[u], a[u]);
      } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
          const bool in = idx[u] + k < numel;
          g[u][k] = in ? to_f32(src[idx[u] + k]) : 0.0f;
          if constexpr (READ_ACC) a[u][k] = in ? acc[dst0 + idx[u] + k] : 0.0f;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
#pragma unroll
      for (int k = 0; k < VEC; ++k) {
        float v = g[u][k];
        if constexpr (READ_ACC) v += a[u][k];
        if constexpr (WRITE_WIRE) v *= scale;
        g[u][k] = v;
      }
      if (idx[u] + VEC <= numel) {
        if constexpr (WRITE_WIRE && PUSH) store_vec<W, VEC>(push_ptr<W>(D, dst0 + idx[u]), g[u]);
        else if constexpr (WRITE_WIRE) store_vec<W, VEC>(wire + dst0 + idx[u], g[u]);
        else store_vec<float, 4>(acc + dst0 + idx[u], g[u]);
      } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
          if (idx[u] + k < numel) {
            if constexpr (WRITE_WIRE && PUSH) *push_ptr<W>(D, dst0 + idx[u] + k) = from_f32<W>(g[u][k]);
            else if constexpr (WRITE_WIRE) wire[dst0 + idx[u] + k] = from_f32<W>(g[u][k]);
            else acc[dst0 + idx[u] + k] = g[u][k];
          }
        }
      }
    }
  }
}

template <typename W>
__global__ void __launch_bounds__(kThreads)
scale_cast_kernel(const float* __restrict__ acc, W* __restrict__ wire, long long n, float scale) {
  const long long nvec = n >> 2;
  const long long stride = (long long)gridDim.x * kThreads;
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < nvec; i += stride) {
    float v[4];
    load_vec<float, 4, Cache::Stream>(acc + i * 4, v);
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] *= scale;
    store_vec<W, 4>(wire + i * 4, v);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long long i = (nvec << 2) + threadIdx.x;
    wire[i] = from_f32<W>(acc[i] * scale);
  }
}

const PushDst kNoPush = {{nullptr}, 0, 0, 0, 0, 0};

template <typename G, typename W, bool RA, bool WW, int kUnroll>
int launch_flatten_u(const dgs_seg* segs, int n_segs, float* acc, void* wire, float scale,
                   cudaStream_t st, const PushDst* push = nullptr) {
  constexpr int TILE = FlatCfg<G, W, RA, WW, kUnroll>::TILE;
  for (int s0 = 0; s0 < n_segs; s0 += kMaxSegs) {
    FlatParams P;
    const int n = (n_segs - s0 < kMaxSegs) ? n_segs - s0 : kMaxSegs;
    long long tiles = 0;
    int m = 0;
    for (int i = 0; i < n; ++i) {
      const dgs_seg& s = segs[s0 + i];
      if (s.numel == 0) continue;
      P.seg[m].src = s.src;
      P.seg[m].dst_off = s.dst_off;
      P.seg[m].numel = s.numel;
      P.seg[m].tile_start = tiles;
      P.seg[m].vec_ok = aligned16(s.src) ? 1 : 0;
      P.seg[m].pad_ = 0;
      tiles += (s.numel + TILE - 1) / TILE;
      ++m;
    }
    if (m == 0) continue;
    P.n = m;
    P.tiles = tiles;
    const long long cap = (long long)sm_count() * grid_mult();
    const int grid = (int)(tiles < cap ? tiles : cap);
    if constexpr (WW) {
      if (push != nullptr) {
        PushDst D = *push;
        D.per = (tiles + D.n - 1) / D.n;
        const long long visit = D.per * D.n;
        const int pgrid = (int)(visit < cap ? visit : cap);
        flatten_kernel<G, W, RA, WW, kUnroll, true><<<pgrid, kThreads, 0, st>>>(P, acc, static_cast<W*>(wire), scale, D);
        DGS_LAUNCHED();
        continue;
      }
    }
    flatten_kernel<G, W, RA, WW, kUnroll, false><<<grid, kThreads, 0, st>>>(P, acc, static_cast<W*>(wire), scale, kNoPush);
    DGS_LAUNCHED();
  }
  return DGS_OK;
}

inline int flat_unroll() {
  static const int v = [] {
    const char* e = getenv("DGS_FLAT_UNROLL");
    return e ? atoi(e) : 0;
  }();
  return v;
}

template <typename G, typename W, bool RA, bool WW>
int launch_flatten(const dgs_seg* segs, int n_segs, float* acc, void* wire, float scale, cudaStream_t st,
                   const PushDst* push = nullptr) {
  // 2 vectors per thread per array measured best on B200 for every dtype mix (gpurun_out/sweep_unroll*.txt:
  // bf16->bf16 0.78 -> 1.00 of the copy peak against 4); DGS_FLAT_UNROLL=1|4 re-runs the sweep.
  const int u = flat_unroll();
  if (u == 4) return launch_flatten_u<G, W, RA, WW, 4>(segs, n_segs, acc, wire, scale, st, push);
  if (u == 1) return launch_flatten_u<G, W, RA, WW, 1>(segs, n_segs, acc, wire, scale, st, push);
  return launch_flatten_u<G, W, RA, WW, 2>(segs, n_segs, acc, wire, scale, st, push);
}

template <typename G, typename W>
int dispatch_mode(int mode, const dgs_seg* segs, int n, float* acc, void* wire, float scale,
                  cudaStream_t st, const PushDst* push = nullptr) {
  switch (mode) {
    case 0: return launch_flatten<G, W, false, false>(segs, n, acc, wire, scale, st);
    case DGS_READ_ACC: return launch_flatten<G, W, true, false>(segs, n, acc, wire, scale, st);
    case DGS_WRITE_WIRE: return launch_flatten<G, W, false, true>(segs, n, acc, wire, scale, st, push);
    default: return launch_flatten<G, W, true, true>(segs, n, acc, wire, scale, st, push);
  }
}

template <typename G>
int dispatch_wire(int wire_dtype, int mode, const dgs_seg* segs, int n, float* acc, void* wire,
                  float scale, cudaStream_t st, const PushDst* push = nullptr) {
  switch (wire_dtype) {
    case DGS_F32: return dispatch_mode<G, float>(mode, segs, n, acc, wire, scale, st, push);
    case DGS_BF16: return dispatch_mode<G, __nv_bfloat16>(mode, segs, n, acc, wire, scale, st, push);
    case DGS_F16: return dispatch_mode<G, __half>(mode, segs, n, acc, wire, scale, st, push);
    default: return fail(DGS_EINVAL, "dgs_flatten_accumulate: bad wire_dtype %d", wire_dtype);
  }
}

int flatten_entry(const char* who, const dgs_seg* segs, int n_segs, int grad_dtype, float* acc, void* wire, int wire_dtype,
                  int mode, float scale, const PushDst* push, cudaStream_t st) {
  DGS_CHECK_ARG(n_segs >= 0, "%s: n_segs %d < 0", who, n_segs);
  if (n_segs == 0) return DGS_OK;
  DGS_CHECK_ARG(segs != nullptr, "%s: segs is NULL", who);
  DGS_CHECK_ARG((mode & ~(DGS_READ_ACC | DGS_WRITE_WIRE)) == 0, "%s: bad mode %d", who, mode);
  const bool ww = (mode & DGS_WRITE_WIRE) != 0, ra = (mode & DGS_READ_ACC) != 0;
  DGS_CHECK_ARG(!ww || wire != nullptr, "%s: DGS_WRITE_WIRE needs a wire buffer", who);
  DGS_CHECK_ARG((!ra && ww) || acc != nullptr, "%s: this mode needs the accumulator", who);
  DGS_CHECK_ARG(acc == nullptr || aligned16(acc), "%s: acc is not 16-byte aligned", who);
  DGS_CHECK_ARG(wire == nullptr || aligned16(wire), "%s: wire is not 16-byte aligned", who);
  for (int i = 0; i < n_segs; ++i) {
    DGS_CHECK_ARG(segs[i].numel >= 0 && segs[i].dst_off >= 0, "%s: seg %d negative size/offset", who, i);
    DGS_CHECK_ARG(segs[i].numel == 0 || segs[i].src != nullptr, "%s: seg %d src is NULL", who, i);
    DGS_CHECK_ARG((segs[i].dst_off & 7) == 0, "%s: seg %d dst_off %lld is not a multiple of 8", who, i,
                  (long long)segs[i].dst_off);
    if (push != nullptr)
      DGS_CHECK_ARG(segs[i].dst_off >= push->begin && segs[i].dst_off + segs[i].numel <= push->begin + push->n * push->span,
                    "%s: seg %d [%lld, +%lld) lies outside the bucket [%lld, +%lld)", who, i, (long long)segs[i].dst_off,
                    (long long)segs[i].numel, (long long)push->begin, (long long)(push->n * push->span));
  }
  if (!ww) wire_dtype = DGS_F32;  // unused
  switch (grad_dtype) {
    case DGS_F32: return dispatch_wire<float>(wire_dtype, mode, segs, n_segs, acc, wire, scale, st, push);
    case DGS_BF16: return dispatch_wire<__nv_bfloat16>(wire_dtype, mode, segs, n_segs, acc, wire, scale, st, push);
    case DGS_F16: return dispatch_wire<__half>(wire_dtype, mode, segs, n_segs, acc, wire, scale, st, push);
    default: return fail(DGS_EINVAL, "%s: bad grad_dtype %d", who, grad_dtype);
  }
}

}  // namespace
}  // namespace dgs

extern "C" int dgs_flatten_accumulate(const dgs_seg* segs, int n_segs, int grad_dtype, float* acc,
                                      void* wire, int wire_dtype, int mode, float scale,
                                      dgs_stream_t stream) {
  return dgs::flatten_entry("dgs_flatten_accumulate", segs, n_segs, grad_dtype, acc, wire, wire_dtype, mode, scale, nullptr,
                            static_cast<cudaStream_t>(stream));
}

extern "C" int dgs_flatten_push(const dgs_seg* segs, int n_segs, int grad_dtype, float* acc, void* const* wire_peers,
                                int wire_dtype, int mode, float scale, int64_t bucket_begin, int64_t span, int nranks,
                                int rank, dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(nranks >= 2 && nranks <= DGS_PEER_MAX && rank >= 0 && rank < nranks, "dgs_flatten_push: rank %d of %d (at most %d ranks)",
                rank, nranks, DGS_PEER_MAX);
  DGS_CHECK_ARG((mode & DGS_WRITE_WIRE) != 0, "dgs_flatten_push: only the modes that write the wire push (use dgs_flatten_accumulate)");
  DGS_CHECK_ARG(wire_peers != nullptr, "dgs_flatten_push: wire_peers is NULL");
  DGS_CHECK_ARG(bucket_begin >= 0 && (bucket_begin & 7) == 0 && span >= 0 && (span & 7) == 0,
                "dgs_flatten_push: bucket_begin %lld and span %lld must be multiples of 8 elements", (long long)bucket_begin, (long long)span);
  PushDst D;
  for (int r = 0; r < DGS_PEER_MAX; ++r) D.wire[r] = nullptr;
  for (int r = 0; r < nranks; ++r) {
    DGS_CHECK_ARG(wire_peers[r] != nullptr && aligned16(wire_peers[r]), "dgs_flatten_push: wire of rank %d is NULL or misaligned", r);
    D.wire[r] = wire_peers[r];
  }
  D.begin = bucket_begin;
  D.span = span;
  D.per = 0;
  D.n = nranks;
  D.rank = rank;
  return flatten_entry("dgs_flatten_push", segs, n_segs, grad_dtype, acc, wire_peers[rank], wire_dtype, mode, scale, &D,
                       static_cast<cudaStream_t>(stream));
}

extern "C" int dgs_scale_cast(const float* acc, void* wire, int wire_dtype, int64_t n, float scale,
                              dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(n >= 0, "dgs_scale_cast: n < 0");
  if (n == 0) return DGS_OK;
  DGS_CHECK_ARG(acc != nullptr && wire != nullptr, "dgs_scale_cast: NULL buffer");
  DGS_CHECK_ARG(aligned16(acc) && aligned16(wire), "dgs_scale_cast: buffers must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long nvec = (n + 3) / 4;
  const long long want = (nvec + kThreads - 1) / kThreads;
  const long long cap = (long long)sm_count() * grid_mult();
  const int grid = (int)(want < cap ? want : cap);
  switch (wire_dtype) {
    case DGS_F32: scale_cast_kernel<float><<<grid, kThreads, 0, st>>>(acc, (float*)wire, n, scale); break;
    case DGS_BF16: scale_cast_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>(acc, (__nv_bfloat16*)wire, n, scale); break;
    case DGS_F16: scale_cast_kernel<__half><<<grid, kThreads, 0, st>>>(acc, (__half*)wire, n, scale); break;
    default: return fail(DGS_EINVAL, "dgs_scale_cast: bad wire_dtype %d", wire_dtype);
  }
  DGS_LAUNCHED();
  return DGS_OK;
}
