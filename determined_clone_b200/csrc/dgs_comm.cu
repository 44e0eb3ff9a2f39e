// Communicator: NCCL over NVLink 5 / NVSwitch on a high-priority side stream.
//
// Takes over dist.init_process_group (reference harness/determined/pytorch/_pytorch_trial.py:296-300),
// the bucket allreduce and broadcasts torch DDP issues for the wrapped model
// (harness/determined/pytorch/_pytorch_context.py:296), hvd.broadcast_parameters/optimizer_state
// (_pytorch_trial.py:617-620) and the tensor payload of DistributedContext.gather/allgather
// (harness/determined/core/_distributed.py:287-356).
//
// libnccl.so.2 is resolved with dlopen so the library loads on machines without NCCL (CPU test
// boxes) and shares the copy torch already mapped into the process.
#include "dgs_common.cuh"

#include <dlfcn.h>
#include <nccl.h>

#include <cstdio>
#include <cstring>
#include <mutex>

struct dgs_comm {
  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1, device = 0;
  cudaStream_t side = nullptr;       // collectives
  cudaStream_t aux = nullptr;        // kernels that consume a collective's result (norm pass 1), so that they
                                     // do not delay the next collective on `side`
  cudaEvent_t ev_compute = nullptr;  // "everything enqueued so far on the compute stream"
  cudaEvent_t ev_side = nullptr;     // "everything issued so far on the side stream"
  cudaEvent_t ev_aux = nullptr;      // "everything issued so far on the aux stream"
  bool pending = false;
  bool aux_pending = false;
};

namespace dgs {
namespace {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*ReduceScatter)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  const char* (*GetLastError)(ncclComm_t) = nullptr;
  char why[256] = {0};
};

NcclApi api;

NcclApi* nccl() {
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle) break;
    }
    if (!api.handle) {
      snprintf(api.why, sizeof(api.why), "dlopen(libnccl.so.2) failed: %s", dlerror());
      return;
    }
#define DGS_SYM(field, name)                                                    \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, name));   \
  if (!api.field) {                                                             \
    snprintf(api.why, sizeof(api.why), "libnccl.so.2 has no symbol %s", name);  \
    api.handle = nullptr;                                                       \
    return;                                                                     \
  }
    DGS_SYM(GetVersion, "ncclGetVersion")
    DGS_SYM(GetUniqueId, "ncclGetUniqueId")
    DGS_SYM(CommInitRank, "ncclCommInitRank")
    DGS_SYM(CommDestroy, "ncclCommDestroy")
    DGS_SYM(AllReduce, "ncclAllReduce")
    DGS_SYM(Broadcast, "ncclBroadcast")
    DGS_SYM(AllGather, "ncclAllGather")
    DGS_SYM(ReduceScatter, "ncclReduceScatter")
    DGS_SYM(GetErrorString, "ncclGetErrorString")
    DGS_SYM(GetLastError, "ncclGetLastError")
#undef DGS_SYM
  });
  return api.handle ? &api : nullptr;
}

int no_nccl() { return fail(DGS_ENCCL, "NCCL is unavailable in this process: %s", api.why); }

#define DGS_NCCL(api, comm, expr)                                                              \
  do {                                                                                         \
    ncclResult_t r__ = (expr);                                                                 \
    if (r__ != ncclSuccess)                                                                    \
      return ::dgs::fail(DGS_ENCCL, "%s failed: %s (%s)", #expr, (api)->GetErrorString(r__), \
                         (api)->GetLastError(comm));                                           \
  } while (0)

bool nccl_dtype(int dt, ncclDataType_t* out) {
  switch (dt) {
    case DGS_F32: *out = ncclFloat32; return true;
    case DGS_BF16: *out = ncclBfloat16; return true;
    case DGS_F16: *out = ncclFloat16; return true;
    case DGS_F64: *out = ncclFloat64; return true;
    case DGS_I64: *out = ncclInt64; return true;
    default: return false;
  }
}

// side stream waits for everything enqueued so far on `compute`
int fence_in(dgs_comm* c, cudaStream_t compute) {
  DGS_CUDA(cudaEventRecord(c->ev_compute, compute));
  DGS_CUDA(cudaStreamWaitEvent(c->side, c->ev_compute, 0));
  return DGS_OK;
}
int mark_out(dgs_comm* c) {
  DGS_CUDA(cudaEventRecord(c->ev_side, c->side));
  c->pending = true;
  return DGS_OK;
}

}  // namespace
}  // namespace dgs

extern "C" int dgs_nccl_version(int* version) {
  using namespace dgs;
  DGS_CHECK_ARG(version != nullptr, "dgs_nccl_version: NULL argument");
  NcclApi* api = nccl();
  if (!api) return no_nccl();
  DGS_NCCL(api, nullptr, api->GetVersion(version));
  return DGS_OK;
}

extern "C" int dgs_comm_unique_id(void* out128) {
  using namespace dgs;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  DGS_CHECK_ARG(out128 != nullptr, "dgs_comm_unique_id: NULL argument");
  NcclApi* api = nccl();
  if (!api) return no_nccl();
  ncclUniqueId id;
  DGS_NCCL(api, nullptr, api->GetUniqueId(&id));
  memcpy(out128, &id, sizeof(id));
  return DGS_OK;
}

extern "C" int dgs_comm_init(const void* id128, int rank, int nranks, int device, dgs_comm** out) {
  using namespace dgs;
  DGS_CHECK_ARG(id128 && out, "dgs_comm_init: NULL argument");
  DGS_CHECK_ARG(nranks >= 1 && rank >= 0 && rank < nranks, "dgs_comm_init: rank %d of %d", rank, nranks);
  NcclApi* api = nccl();
  if (!api) return no_nccl();
  DGS_CUDA(cudaSetDevice(device));
  dgs_comm* c = new dgs_comm();
  c->rank = rank;
  c->nranks = nranks;
  c->device = device;
  int lo = 0, hi = 0;
  DGS_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));  // hi = numerically lowest = highest priority
  DGS_CUDA(cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, hi));
  DGS_CUDA(cudaStreamCreateWithPriority(&c->aux, cudaStreamNonBlocking, hi));
  DGS_CUDA(cudaEventCreateWithFlags(&c->ev_compute, cudaEventDisableTiming));
  DGS_CUDA(cudaEventCreateWithFlags(&c->ev_side, cudaEventDisableTiming));
  DGS_CUDA(cudaEventCreateWithFlags(&c->ev_aux, cudaEventDisableTiming));
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  DGS_NCCL(api, nullptr, api->CommInitRank(&c->comm, nranks, id, rank));
  *out = c;
  return DGS_OK;
}

extern "C" int dgs_comm_destroy(dgs_comm* c) {
  using namespace dgs;
  if (!c) return DGS_OK;
  NcclApi* api = nccl();
  if (c->side) (void)cudaStreamSynchronize(c->side);
  if (c->aux) (void)cudaStreamSynchronize(c->aux);
  if (api && c->comm) (void)api->CommDestroy(c->comm);
  if (c->ev_compute) (void)cudaEventDestroy(c->ev_compute);
  if (c->ev_side) (void)cudaEventDestroy(c->ev_side);
  if (c->ev_aux) (void)cudaEventDestroy(c->ev_aux);
  if (c->side) (void)cudaStreamDestroy(c->side);
  if (c->aux) (void)cudaStreamDestroy(c->aux);
  delete c;
  return DGS_OK;
}

extern "C" int dgs_comm_rank(const dgs_comm* c, int* rank, int* nranks) {
  using namespace dgs;
  DGS_CHECK_ARG(c && rank && nranks, "dgs_comm_rank: NULL argument");
  *rank = c->rank;
  *nranks = c->nranks;
  return DGS_OK;
}

extern "C" int dgs_comm_allreduce_async(dgs_comm* c, void* buf, int64_t count, int dtype, dgs_stream_t compute) {
  using namespace dgs;
  DGS_CHECK_ARG(c && c->comm, "dgs_comm_allreduce_async: NULL communicator");
  DGS_CHECK_ARG(count >= 0, "dgs_comm_allreduce_async: count < 0");
  if (count == 0) return DGS_OK;
  DGS_CHECK_ARG(buf != nullptr, "dgs_comm_allreduce_async: NULL buffer");
  ncclDataType_t dt;
  DGS_CHECK_ARG(nccl_dtype(dtype, &dt), "dgs_comm_allreduce_async: bad dtype %d", dtype);
  NcclApi* api = nccl();
  if (int rc = fence_in(c, static_cast<cudaStream_t>(compute))) return rc;
  DGS_NCCL(api, c->comm, api->AllReduce(buf, buf, (size_t)count, dt, ncclSum, c->comm, c->side));
  return mark_out(c);
}

extern "C" int dgs_comm_reduce_scatter_async(dgs_comm* c, void* buf, int64_t count_per_rank, int dtype,
                                             dgs_stream_t compute) {
  using namespace dgs;
  DGS_CHECK_ARG(c && c->comm, "dgs_comm_reduce_scatter_async: NULL communicator");
  DGS_CHECK_ARG(count_per_rank >= 0, "dgs_comm_reduce_scatter_async: count_per_rank < 0");
  if (count_per_rank == 0) return DGS_OK;
  DGS_CHECK_ARG(buf != nullptr, "dgs_comm_reduce_scatter_async: NULL buffer");
  ncclDataType_t dt;
  DGS_CHECK_ARG(nccl_dtype(dtype, &dt), "dgs_comm_reduce_scatter_async: bad dtype %d", dtype);
  NcclApi* api = nccl();
  if (int rc = fence_in(c, static_cast<cudaStream_t>(compute))) return rc;
  char* mine = static_cast<char*>(buf) + (size_t)c->rank * (size_t)count_per_rank * (size_t)dtype_size(dtype);
  DGS_NCCL(api, c->comm, api->ReduceScatter(buf, mine, (size_t)count_per_rank, dt, ncclSum, c->comm, c->side));
  return mark_out(c);
}

extern "C" int dgs_comm_broadcast_async(dgs_comm* c, void* buf, int64_t bytes, int root, dgs_stream_t compute) {
  using namespace dgs;
  DGS_CHECK_ARG(c && c->comm, "dgs_comm_broadcast_async: NULL communicator");
  DGS_CHECK_ARG(bytes >= 0 && root >= 0 && root < c->nranks, "dgs_comm_broadcast_async: bad bytes/root");
  if (bytes == 0) return DGS_OK;
  DGS_CHECK_ARG(buf != nullptr, "dgs_comm_broadcast_async: NULL buffer");
  NcclApi* api = nccl();
  if (int rc = fence_in(c, static_cast<cudaStream_t>(compute))) return rc;
  DGS_NCCL(api, c->comm, api->Broadcast(buf, buf, (size_t)bytes, ncclUint8, root, c->comm, c->side));
  return mark_out(c);
}

extern "C" int dgs_comm_allgather_async(dgs_comm* c, const void* send, void* recv, int64_t bytes_per_rank,
                                        dgs_stream_t compute) {
  using namespace dgs;
  DGS_CHECK_ARG(c && c->comm, "dgs_comm_allgather_async: NULL communicator");
  DGS_CHECK_ARG(bytes_per_rank >= 0, "dgs_comm_allgather_async: bytes_per_rank < 0");
  if (bytes_per_rank == 0) return DGS_OK;
  DGS_CHECK_ARG(send && recv, "dgs_comm_allgather_async: NULL buffer");
  NcclApi* api = nccl();
  if (int rc = fence_in(c, static_cast<cudaStream_t>(compute))) return rc;
  DGS_NCCL(api, c->comm, api->AllGather(send, recv, (size_t)bytes_per_rank, ncclUint8, c->comm, c->side));
  return mark_out(c);
}

extern "C" int dgs_comm_l2norm_partial_async(dgs_comm* c, const void* x, int dtype, int64_t n, float* partials) {
  using namespace dgs;
  DGS_CHECK_ARG(c != nullptr, "dgs_comm_l2norm_partial_async: NULL communicator");
  // aux waits for the collectives issued so far (ev_side was recorded behind the last one), then reads
  DGS_CUDA(cudaStreamWaitEvent(c->aux, c->ev_side, 0));  // no-op if nothing was ever issued
  if (int rc = dgs_l2norm_partial(x, dtype, n, partials, c->aux)) return rc;
  DGS_CUDA(cudaEventRecord(c->ev_aux, c->aux));
  c->aux_pending = true;
  return DGS_OK;
}

extern "C" int dgs_comm_wait(dgs_comm* c, dgs_stream_t compute) {
  using namespace dgs;
  DGS_CHECK_ARG(c != nullptr, "dgs_comm_wait: NULL communicator");
  if (c->pending) {
    DGS_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(compute), c->ev_side, 0));
    c->pending = false;
  }
  if (c->aux_pending) {
    DGS_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(compute), c->ev_aux, 0));
    c->aux_pending = false;
  }
  return DGS_OK;
}
