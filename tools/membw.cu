// Standalone HBM bandwidth probe for B200: what read/write mix, cache hint, grid size and unroll does a
// streaming kernel need to reach the copy roofline?  Informs the launch shapes in csrc/.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o /tmp/membw tools/membw.cu && /tmp/membw
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                     \
  do {                                                                            \
    cudaError_t e = (x);                                                          \
    if (e != cudaSuccess) {                                                       \
      printf("%s: %s\n", #x, cudaGetErrorString(e));                              \
      exit(1);                                                                    \
    }                                                                             \
  } while (0)

enum { ST_DEFAULT = 0, ST_CS = 1, ST_WT = 2 };
enum { LD_DEFAULT = 0, LD_CS = 1, LD_NC_NA = 2 };

template <int LD>
__device__ __forceinline__ float4 ld(const float4* p) {
  if constexpr (LD == LD_CS) return __ldcs(p);
  if constexpr (LD == LD_NC_NA) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
  }
  return *p;
}
template <int ST>
__device__ __forceinline__ void st(float4* p, float4 v) {
  if constexpr (ST == ST_CS) __stcs(p, v);
  else if constexpr (ST == ST_WT) __stwt(p, v);
  else *p = v;
}

// NR read streams, NW write streams, tile-strided persistent loop
template <int NR, int NW, int UNROLL, int LD, int ST, int THREADS>
__global__ void __launch_bounds__(THREADS) stream_kernel(const float4* __restrict__ a, const float4* __restrict__ b,
                                                     float4* __restrict__ c, float4* __restrict__ d, long long nvec) {
  constexpr int TILE = THREADS * UNROLL;
  const long long tiles = nvec / TILE;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    float4 x[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long i = t * TILE + u * THREADS + threadIdx.x;
      x[u] = make_float4(1.f, 2.f, 3.f, 4.f);
      if constexpr (NR >= 1) x[u] = ld<LD>(a + i);
      if constexpr (NR >= 2) {
        float4 y = ld<LD>(b + i);
        x[u].x += y.x; x[u].y += y.y; x[u].z += y.z; x[u].w += y.w;
      }
      if constexpr (NR >= 3) {
        float4 y = ld<LD>(c + i);
        x[u].x += y.x; x[u].y += y.y; x[u].z += y.z; x[u].w += y.w;
      }
      if constexpr (NR >= 4) {
        float4 y = ld<LD>(d + i);
        x[u].x += y.x; x[u].y += y.y; x[u].z += y.z; x[u].w += y.w;
      }
    }
    if constexpr (NW == 0) {
      float s = 0.f;
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) s += x[u].x + x[u].y + x[u].z + x[u].w;
      if (s == 123.456f) c[0] = x[0];
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long i = t * TILE + u * THREADS + threadIdx.x;
      if constexpr (NW >= 1) st<ST>(c + i, x[u]);
      if constexpr (NW >= 2) st<ST>(d + i, x[u]);
      if constexpr (NW >= 3) st<ST>(const_cast<float4*>(b) + i, x[u]);
    }
  }
}

template <int NR, int NW, int UNROLL, int LD, int ST, int THREADS = 256>
void run(const char* name, int blocks_per_sm, float4* a, float4* b, float4* c, float4* d, long long nvec) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  const int grid = 148 * blocks_per_sm;
  float best = 1e9f, sum = 0.f;
  const int iters = 12;
  for (int it = 0; it < iters + 3; ++it) {
    CK(cudaEventRecord(e0));
    stream_kernel<NR, NW, UNROLL, LD, ST, THREADS><<<grid, THREADS>>>(a, b, c, d, nvec);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (it >= 3) {
      sum += ms;
      if (ms < best) best = ms;
    }
  }
  const double bytes = (double)(NR + NW) * nvec * 16;
  printf("%-34s R%d W%d thr %d unroll %d grid 148x%-2d ld %d st %d : avg %.4f ms  %.0f GB/s   best %.0f GB/s\n", name, NR, NW, THREADS, UNROLL,
         blocks_per_sm, LD, ST, sum / iters, bytes / (sum / iters * 1e-3) / 1e9, bytes / (best * 1e-3) / 1e9);
}

int main(int argc, char** argv) {
  const long long mb = argc > 1 ? atoll(argv[1]) : 1024;
  const long long nvec = mb * (1 << 20) / 16;
  float4 *a, *b, *c, *d;
  CK(cudaMalloc(&a, nvec * 16));
  CK(cudaMalloc(&b, nvec * 16));
  CK(cudaMalloc(&c, nvec * 16));
  CK(cudaMalloc(&d, nvec * 16));
  CK(cudaMemset(a, 0, nvec * 16));
  CK(cudaMemset(b, 0, nvec * 16));
  printf("array = %lld MiB\n", mb);
  // memcpy yardstick
  {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float sum = 0;
    for (int it = 0; it < 13; ++it) {
      CK(cudaEventRecord(e0));
      CK(cudaMemcpyAsync(c, a, nvec * 16, cudaMemcpyDeviceToDevice));
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (it >= 3) sum += ms;
    }
    printf("cudaMemcpyAsync D2D: %.4f ms %.0f GB/s\n", sum / 10, 2.0 * nvec * 16 / (sum / 10 * 1e-3) / 1e9);
  }
  run<1, 1, 4, LD_CS, ST_DEFAULT, 256>("copy", 1, a, b, c, d, nvec);
  run<1, 1, 4, LD_CS, ST_DEFAULT, 256>("copy", 2, a, b, c, d, nvec);
  run<1, 1, 4, LD_CS, ST_DEFAULT, 256>("copy", 3, a, b, c, d, nvec);
  run<1, 1, 8, LD_CS, ST_DEFAULT, 256>("copy", 1, a, b, c, d, nvec);
  run<1, 1, 8, LD_CS, ST_DEFAULT, 256>("copy", 2, a, b, c, d, nvec);
  run<1, 1, 4, LD_CS, ST_DEFAULT, 512>("copy", 1, a, b, c, d, nvec);
  run<1, 1, 4, LD_CS, ST_DEFAULT, 512>("copy", 2, a, b, c, d, nvec);
  run<1, 1, 4, LD_CS, ST_DEFAULT, 1024>("copy", 1, a, b, c, d, nvec);
  run<1, 1, 2, LD_CS, ST_DEFAULT, 1024>("copy", 1, a, b, c, d, nvec);
  run<1, 1, 2, LD_CS, ST_DEFAULT, 1024>("copy", 2, a, b, c, d, nvec);
  run<1, 1, 8, LD_CS, ST_DEFAULT, 1024>("copy", 1, a, b, c, d, nvec);
  run<2, 1, 4, LD_CS, ST_DEFAULT, 256>("add 2R1W", 2, a, b, c, d, nvec);
  run<2, 1, 4, LD_CS, ST_DEFAULT, 256>("add 2R1W", 4, a, b, c, d, nvec);
  run<2, 1, 4, LD_CS, ST_DEFAULT, 512>("add 2R1W", 2, a, b, c, d, nvec);
  run<2, 1, 2, LD_CS, ST_DEFAULT, 1024>("add 2R1W", 1, a, b, c, d, nvec);
  run<4, 3, 2, LD_DEFAULT, ST_DEFAULT, 256>("adam 4R3W", 1, a, b, c, d, nvec);
  run<4, 3, 2, LD_DEFAULT, ST_DEFAULT, 256>("adam 4R3W", 2, a, b, c, d, nvec);
  run<4, 3, 2, LD_DEFAULT, ST_DEFAULT, 256>("adam 4R3W", 4, a, b, c, d, nvec);
  run<4, 3, 2, LD_DEFAULT, ST_DEFAULT, 256>("adam 4R3W", 8, a, b, c, d, nvec);
  run<4, 3, 1, LD_DEFAULT, ST_DEFAULT, 256>("adam 4R3W", 4, a, b, c, d, nvec);
  run<4, 3, 1, LD_DEFAULT, ST_DEFAULT, 256>("adam 4R3W", 8, a, b, c, d, nvec);
  run<4, 3, 2, LD_DEFAULT, ST_DEFAULT, 512>("adam 4R3W", 1, a, b, c, d, nvec);
  run<4, 3, 2, LD_DEFAULT, ST_DEFAULT, 512>("adam 4R3W", 2, a, b, c, d, nvec);
  run<4, 3, 1, LD_DEFAULT, ST_DEFAULT, 1024>("adam 4R3W", 1, a, b, c, d, nvec);
  run<4, 3, 2, LD_DEFAULT, ST_DEFAULT, 1024>("adam 4R3W", 1, a, b, c, d, nvec);
  run<4, 3, 2, LD_DEFAULT, ST_DEFAULT, 256>("adam 4R3W", 64, a, b, c, d, nvec);
  run<3, 2, 2, LD_DEFAULT, ST_DEFAULT, 256>("sgd 3R2W", 2, a, b, c, d, nvec);
  run<3, 2, 2, LD_DEFAULT, ST_DEFAULT, 256>("sgd 3R2W", 8, a, b, c, d, nvec);
  run<3, 2, 2, LD_DEFAULT, ST_DEFAULT, 512>("sgd 3R2W", 2, a, b, c, d, nvec);
  return 0;
}
