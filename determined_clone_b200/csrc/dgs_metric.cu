// Metric reducer kernels: device-side replacement for the per-batch `.cpu()` + numpy averaging of
// training metrics.
//
// Reference: harness/determined/pytorch/_pytorch_trial.py:920-927 (per-batch D2H),
// harness/determined/pytorch/_metric_utils.py:13-60 (_process_combined_metrics_and_batches,
// _average_training_metrics), harness/determined/util.py:171-196 (make_metrics batch mean) and
// harness/determined/pytorch/_reducer.py:28-46 (_simple_reduce_metrics).
//
// Sizes are tiny (B x M scalars); the point is removing the host sync from every batch, not bandwidth.
// Sums run in rank order / batch order in double, counts are int64 and exact.
#include "dgs_common.cuh"

#include <math_constants.h>

namespace dgs {
namespace {

struct PackParams {
  const void* ptr[DGS_MAX_METRICS];
  int dtype[DGS_MAX_METRICS];
  int m;
};

__global__ void metric_pack_kernel(const __grid_constant__ PackParams P, double* __restrict__ row,
                                   uint8_t* __restrict__ valid) {
  const int j = threadIdx.x;
  if (j >= P.m) return;
  const void* p = P.ptr[j];
  double v = 0.0;
  if (p != nullptr) {
    switch (P.dtype[j]) {
      case DGS_F32: v = (double)*static_cast<const float*>(p); break;
      case DGS_BF16: v = (double)__bfloat162float(*static_cast<const __nv_bfloat16*>(p)); break;
      case DGS_F16: v = (double)__half2float(*static_cast<const __half*>(p)); break;
      case DGS_F64: v = *static_cast<const double*>(p); break;
      case DGS_I64: v = (double)*static_cast<const long long*>(p); break;
    }
  }
  row[j] = v;
  valid[j] = p != nullptr;
}

// one thread per (batch, metric): mean over contributing ranks, in rank order.  Rank r's values, validity
// bytes and batch count start r * stride_* bytes after the base pointers (separate arrays or packed records).
__global__ void metric_rank_mean_kernel(const char* __restrict__ gathered, const char* __restrict__ valid,
                                        const char* __restrict__ num_batches, long long stride_vals,
                                        long long stride_valid, long long stride_nb, int n_ranks, int b, int m,
                                        double* __restrict__ per_batch, uint8_t* __restrict__ per_batch_valid) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= b * m) return;
  double s = 0.0;
  long long c = 0;
  for (int r = 0; r < n_ranks; ++r) {
    if (*reinterpret_cast<const long long*>(num_batches + r * stride_nb) == 0) continue;  // rank reported nothing
    if (reinterpret_cast<const uint8_t*>(valid + r * stride_valid)[i]) {
      s += reinterpret_cast<const double*>(gathered + r * stride_vals)[i];
      ++c;
    }
  }
  per_batch[i] = c ? s / (double)c : 0.0;
  per_batch_valid[i] = c != 0;
}

// one block per metric: mean over batches of the valid per-batch values (util.make_metrics)
__global__ void __launch_bounds__(256)
metric_batch_mean_kernel(const double* __restrict__ per_batch, const uint8_t* __restrict__ per_batch_valid, int b,
                         int m, double* __restrict__ avg, long long* __restrict__ count) {
  __shared__ double red[32];
  __shared__ double redc[32];
  const int j = blockIdx.x;
  double s = 0.0, c = 0.0;
  for (int i = threadIdx.x; i < b; i += blockDim.x) {
    if (per_batch_valid[(size_t)i * m + j]) {
      s += per_batch[(size_t)i * m + j];
      c += 1.0;
    }
  }
  s = block_sum(s, red);
  __syncthreads();
  c = block_sum(c, redc);
  if (threadIdx.x == 0) {
    const long long n = (long long)c;
    count[j] = n;
    avg[j] = n ? s / (double)n : CUDART_NAN;
  }
}

// one thread per metric: Reducer.AVG / SUM / MAX / MIN across n entries
__global__ void metric_reduce_kernel(const double* __restrict__ values, const long long* __restrict__ weights,
                                     int n, int m, int op, double* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  double r;
  if (op == DGS_REDUCE_AVG) {
    double s = 0.0, w = 0.0;
    for (int i = 0; i < n; ++i) {
      const double wi = weights ? (double)weights[i] : 1.0;
      s += values[(size_t)i * m + j] * wi;  // np.average: multiply(a, wgt).sum() / wgt.sum()
      w += wi;
    }
    r = s / w;
  } else if (op == DGS_REDUCE_SUM) {
    r = 0.0;
    for (int i = 0; i < n; ++i) r += values[(size_t)i * m + j];
  } else if (op == DGS_REDUCE_MAX) {
    r = values[j];
    for (int i = 1; i < n; ++i) {
      const double x = values[(size_t)i * m + j];
      if (x > r || x != x) r = x;  // np.max propagates NaN
    }
  } else {
    r = values[j];
    for (int i = 1; i < n; ++i) {
      const double x = values[(size_t)i * m + j];
      if (x < r || x != x) r = x;
    }
  }
  out[j] = r;
}

}  // namespace
}  // namespace dgs

extern "C" int dgs_metric_pack(const dgs_scalar_ref* refs, int m, double* row, uint8_t* valid,
                               dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(m >= 0 && m <= DGS_MAX_METRICS, "dgs_metric_pack: m %d outside [0, %d]", m, DGS_MAX_METRICS);
  if (m == 0) return DGS_OK;
  DGS_CHECK_ARG(refs && row && valid, "dgs_metric_pack: NULL argument");
  PackParams P;
  P.m = m;
  for (int j = 0; j < m; ++j) {
    DGS_CHECK_ARG(dtype_size(refs[j].dtype) > 0, "dgs_metric_pack: metric %d has bad dtype %d", j, refs[j].dtype);
    P.ptr[j] = refs[j].ptr;
    P.dtype[j] = refs[j].dtype;
  }
  metric_pack_kernel<<<1, DGS_MAX_METRICS, 0, static_cast<cudaStream_t>(stream)>>>(P, row, valid);
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_metric_average(const double* gathered, const uint8_t* valid, const int64_t* num_batches,
                                  int n_ranks, int b, int m, double* per_batch, uint8_t* per_batch_valid,
                                  double* avg, int64_t* count, dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(n_ranks >= 1 && b >= 0 && m >= 0, "dgs_metric_average: bad shape [%d, %d, %d]", n_ranks, b, m);
  if (m == 0) return DGS_OK;
  DGS_CHECK_ARG(gathered && valid && num_batches && per_batch && per_batch_valid && avg && count,
                "dgs_metric_average: NULL argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (b > 0) {
    const int total = b * m;
    metric_rank_mean_kernel<<<(total + 255) / 256, 256, 0, st>>>(
        reinterpret_cast<const char*>(gathered), reinterpret_cast<const char*>(valid),
        reinterpret_cast<const char*>(num_batches), 8LL * total, 1LL * total, 8LL, n_ranks, b, m, per_batch, per_batch_valid);
    DGS_LAUNCHED();
  }
  metric_batch_mean_kernel<<<m, 256, 0, st>>>(per_batch, per_batch_valid, b, m, avg,
                                              reinterpret_cast<long long*>(count));
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_metric_average_packed(const void* records, int64_t record_bytes, int n_ranks, int b, int m,
                                         double* per_batch, uint8_t* per_batch_valid, double* avg, int64_t* count,
                                         dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(n_ranks >= 1 && b >= 0 && m >= 0, "dgs_metric_average_packed: bad shape [%d, %d, %d]", n_ranks, b, m);
  if (m == 0) return DGS_OK;
  const long long total = (long long)b * m;
  DGS_CHECK_ARG((record_bytes & 7) == 0 && record_bytes >= 9 * total + 8,
                "dgs_metric_average_packed: record_bytes %lld too small or not a multiple of 8", (long long)record_bytes);
  DGS_CHECK_ARG(records && per_batch && per_batch_valid && avg && count, "dgs_metric_average_packed: NULL argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const char* base = static_cast<const char*>(records);
  if (b > 0) {
    metric_rank_mean_kernel<<<(int)((total + 255) / 256), 256, 0, st>>>(
        base, base + 8 * total + 8, base + 8 * total, record_bytes, record_bytes, record_bytes, n_ranks, b, m, per_batch,
        per_batch_valid);
    DGS_LAUNCHED();
  }
  metric_batch_mean_kernel<<<m, 256, 0, st>>>(per_batch, per_batch_valid, b, m, avg, reinterpret_cast<long long*>(count));
  DGS_LAUNCHED();
  return DGS_OK;
}

extern "C" int dgs_metric_reduce(const double* values, const int64_t* weights, int n, int m, int op, double* out,
                                 dgs_stream_t stream) {
  using namespace dgs;
  DGS_CHECK_ARG(n >= 1 && m >= 0, "dgs_metric_reduce: bad shape [%d, %d]", n, m);
  DGS_CHECK_ARG(op >= DGS_REDUCE_AVG && op <= DGS_REDUCE_MIN, "dgs_metric_reduce: bad op %d", op);
  if (m == 0) return DGS_OK;
  DGS_CHECK_ARG(values && out, "dgs_metric_reduce: NULL argument");
  metric_reduce_kernel<<<(m + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      values, reinterpret_cast<const long long*>(weights), n, m, op, out);
  DGS_LAUNCHED();
  return DGS_OK;
}
