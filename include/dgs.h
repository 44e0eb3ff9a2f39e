/*
 * dgs.h -- C ABI of libdgs.so ("determined grad sync"), the B200 (sm_100a) replacement for the
 * per-step data-parallel path of Determined's PyTorch harness.
 *
 * The reference has no FFI for this path: its hot path is Python calling into torch
 * (DistributedDataParallel, torch.optim, clip_grad_norm_) -- see SURVEY.md section 8(b).  Each entry
 * point below therefore cites the reference *call site* (path relative to the reference tree, H/ =
 * harness/determined/) whose work it takes over.
 *
 * Conventions
 *   - plain C: pointers, sizes, scalars.  No torch / C++ types cross this boundary.
 *   - every function returns 0 on success, a DGS_E* code otherwise; dgs_last_error() gives the
 *     thread-local message of the last failure on the calling thread.
 *   - device pointers are owned by the caller.  Unless a function says otherwise the library never
 *     allocates or frees device memory; scratch is passed in by the caller.
 *   - streams are raw cudaStream_t values passed as void*; all work is enqueued asynchronously and
 *     nothing here blocks the host except dgs_comm_init / dgs_comm_destroy.
 *   - "flat buffer": one contiguous device array holding every tensor of one optimizer, each tensor
 *     starting at a multiple of DGS_ALIGN_BYTES bytes (gaps are zero).
 */
#ifndef DGS_H_
#define DGS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DGS_VERSION 100 /* 0.1.0 */
#define DGS_ALIGN_BYTES 128

/* status codes */
#define DGS_OK 0
#define DGS_EINVAL 1       /* bad argument (null pointer, misaligned buffer, unknown dtype, ...) */
#define DGS_ECUDA 2        /* a CUDA runtime call failed */
#define DGS_ENCCL 3        /* an NCCL call failed, or libnccl.so.2 could not be loaded */
#define DGS_EUNSUPPORTED 4 /* valid request this build does not implement */

/* element types */
#define DGS_F32 0
#define DGS_BF16 1
#define DGS_F16 2
#define DGS_F64 3
#define DGS_I64 4

typedef void* dgs_stream_t; /* cudaStream_t */

int dgs_version(void);
const char* dgs_last_error(void);
/* Number of CUDA kernels this library has launched in this process (all threads). */
int64_t dgs_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * 1. flatten / accumulate / scale / cast
 *
 * Replaces: autograd's per-parameter `grad += new` under DDP.no_sync()
 * (H/pytorch/_pytorch_context.py:802-812), DDP's copy of grad/N into its buckets on the boundary
 * backward (:813-816, torch Reducer), and the per-parameter `p.grad.data.div_(A)` loop of
 * _average_gradients (:818-825).
 *
 * One launch handles every gradient tensor of one bucket: tensor i has `numel` elements of
 * `grad_dtype` at `src` and lands at element offset `dst_off` of the bucket.
 *   mode 0                          acc[dst]  = g                      (first micro-batch)
 *   mode DGS_READ_ACC               acc[dst] += g                      (middle micro-batches)
 *   mode DGS_WRITE_WIRE             wire[dst] = cast(g * scale)        (boundary, A == 1)
 *   mode DGS_READ_ACC|DGS_WRITE_WIRE wire[dst] = cast((acc[dst]+g)*scale)  (boundary, A > 1)
 * `acc` is fp32 and may be NULL when it is neither read nor written; `wire` (of `wire_dtype`:
 * F32/BF16/F16) may be NULL when DGS_WRITE_WIRE is not set.  scale is 1/(N*A) for averaged
 * gradients, 1/N when average_aggregated_gradients is false.  `segs` is a HOST array; dst_off must be
 * a multiple of 8 elements, src should be 16-byte aligned (unaligned sources take a scalar path).
 * ---------------------------------------------------------------------------------------------- */
typedef struct dgs_seg {
  const void* src;
  int64_t dst_off;
  int64_t numel;
} dgs_seg;

#define DGS_READ_ACC 1
#define DGS_WRITE_WIRE 2

int dgs_flatten_accumulate(const dgs_seg* segs, int n_segs, int grad_dtype, float* acc, void* wire,
                           int wire_dtype, int mode, float scale, dgs_stream_t stream);

/* wire[i] = cast(acc[i] * scale) for i in [0, n): finishes a bucket whose last gradient never
 * arrived through a hook (same reference lines as above). */
int dgs_scale_cast(const float* acc, void* wire, int wire_dtype, int64_t n, float scale,
                   dgs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * 2. global L2 norm, two passes, no host synchronisation
 *
 * Replaces: the user's clip_grads callable == torch.nn.utils.clip_grad_norm_(params, max_norm)
 * invoked at H/pytorch/_pytorch_context.py:902 (torch: _foreach_norm + vector_norm of the stack).
 *
 * pass 1: dgs_l2norm_partial writes dgs_l2norm_grid(n) per-block sums of squares (fp32) to
 *         `partials` (warp-shuffle + shared-memory block reduction).
 * pass 2: dgs_l2norm_final reduces `n_partials` values in double and writes sqrt(sum) to
 *         out_norm[0] (device fp32), which the step kernels read.
 * Several flat buffers can share one pass 2 by writing their partials side by side.
 * ---------------------------------------------------------------------------------------------- */
#define DGS_NORM_MAX_PARTIALS 9472 /* 148 SMs x 64 */

int dgs_l2norm_grid(int64_t n);
int dgs_l2norm_partial(const void* x, int dtype, int64_t n, float* partials, dgs_stream_t stream);
int dgs_l2norm_final(const float* partials, int n_partials, float* out_norm, dgs_stream_t stream);
/* Pass 2 variants for a norm whose elements are spread over ranks (sharded step): write the SUM of squares
 * instead of its root, all-reduce that one float, then take the root in place. */
int dgs_l2norm_final_sumsq(const float* partials, int n_partials, float* out_sumsq, dgs_stream_t stream);
int dgs_sqrt_inplace(float* x, dgs_stream_t stream);
/* both passes; `partials` must hold dgs_l2norm_grid(n) floats */
int dgs_l2norm(const void* x, int dtype, int64_t n, float* partials, float* out_norm,
               dgs_stream_t stream);

/* The same two-pass norm for callers that cannot allocate device memory at all (a cgo / JNI / ctypes
 * plugin host that owns no device allocator; a trainer whose caching allocator has reserved all of HBM so
 * that a raw cudaMalloc for scratch would fail; step loops under a no-allocation rule): the partials go to
 * a __device__ array built into the library (DGS_NOALLOC_MAX_PARTIALS floats = 16 KiB, resident on every
 * GPU that loads the library until exit) instead of caller scratch.  Accepts n <= DGS_NOALLOC_MAX_ELEMS
 * and refuses larger inputs with DGS_EINVAL.  Calls no cudaMalloc/cudaFree or other allocating routine.
 * The first call on a device resolves the array's device address (do it during set-up, before any stream
 * capture on that device).  One array per device is shared by all calls: calls whose kernels overlap on
 * the GPU race; it keeps the previous call's contents and needs no clearing (pass 1 overwrites every slot
 * pass 2 reads).  This entry point has no counterpart in the reference. */
#define DGS_NOALLOC_MAX_PARTIALS 4096
#define DGS_NOALLOC_MAX_ELEMS (4096LL * 16384LL) /* 2^26 elements */
int dgs_l2norm_noalloc(const void* x, int dtype, int64_t n, float* out_norm, dgs_stream_t stream);

/* g[i] *= min(max_norm / (total_norm[0] + 1e-6), 1) -- the unfused tail of clip_grad_norm_, used
 * only when the optimizer step itself stays in torch (optimizers other than SGD/Adam/AdamW). */
int dgs_clip_scale(void* g, int dtype, int64_t n, const float* total_norm, float max_norm,
                   dgs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * 3. fused clip + optimizer step
 *
 * Replaces: clip_grads(parameters) + optimizer.step() + optimizer.zero_grad() at
 * H/pytorch/_pytorch_context.py:902-924, for torch.optim.AdamW / Adam (amsgrad=False) and
 * torch.optim.SGD.  Formulas follow torch 2.11 torch/optim/adamw.py, adam.py
 * (_single_tensor_adam) and sgd.py (_single_tensor_sgd).
 *
 * param / state / grad are flat buffers with the same element layout.  Hyper-parameters are
 * per range (a run of tensors of one param group with the same step count); elements outside
 * every range are left untouched (parameters that received no gradient).  Range bounds must be
 * multiples of 8 elements (the last `end` may be n).  At most DGS_MAX_RANGES per call.
 * If total_norm != NULL every gradient is first multiplied by
 * min(max_norm / (total_norm[0] + 1e-6), 1)  (torch/nn/utils/clip_grad.py).
 * The accumulator needs no zeroing: the next window's first micro-batch overwrites it.
 * ---------------------------------------------------------------------------------------------- */
#define DGS_MAX_RANGES 24

#define DGS_OPT_MAXIMIZE 1
#define DGS_ADAM_DECOUPLED_WD 2 /* AdamW: p *= 1 - lr*wd ; otherwise Adam: g += wd*p */
#define DGS_SGD_NESTEROV 4
#define DGS_SGD_FIRST_STEP 8 /* momentum buffer does not exist yet: buf = g */

/* Hyper-parameters are doubles, exactly the Python floats torch.optim holds in param_groups; the
 * kernels derive their fp32 constants from them the way torch does (double arithmetic, then one
 * rounding to fp32). */
typedef struct dgs_adam_range {
  int64_t begin, end;
  double lr, beta1, beta2, eps, weight_decay;
  double step_size;             /* lr / (1 - beta1^t) */
  double bias_correction2_sqrt; /* sqrt(1 - beta2^t) */
  int32_t flags;
  int32_t _pad;
} dgs_adam_range;

typedef struct dgs_sgd_range {
  int64_t begin, end;
  double lr, momentum, dampening, weight_decay;
  int32_t flags;
  int32_t _pad;
} dgs_sgd_range;

int dgs_step_adam(void* param, void* exp_avg, void* exp_avg_sq, int state_dtype, const void* grad,
                  int grad_dtype, int64_t n, const dgs_adam_range* ranges, int n_ranges,
                  const float* total_norm, float max_norm, dgs_stream_t stream);

/* momentum_buf may be NULL when every range has momentum == 0 */
int dgs_step_sgd(void* param, void* momentum_buf, int state_dtype, const void* grad, int grad_dtype,
                 int64_t n, const dgs_sgd_range* ranges, int n_ranges, const float* total_norm,
                 float max_norm, dgs_stream_t stream);

/* The same two steps for gradients that still carry a torch.amp.GradScaler loss scale S (AMP, reference
 * H/pytorch/_pytorch_context.py:798-799 scales the loss, :899-911 `scaler.unscale_` + `scaler.step`).  Replaces the
 * separate unscale pass of GradScaler.unscale_ (torch/amp/grad_scaler.py: _amp_foreach_non_finite_check_and_unscale_,
 * one read + one write of every gradient): inv_scale[0] = 1/S is a device scalar, every gradient element is
 * multiplied by it on the fly, and total_norm[0] is expected to be the norm of the SCALED buffer (the clip
 * coefficient is computed from total_norm[0] * inv_scale[0]).  The inf/nan check that decides whether the step
 * runs at all is the caller's: a non-finite total_norm[0] of the scaled buffer means "skip", exactly when some
 * element is inf/nan (or when the fp32 sum of squares overflowed, norm > 1.8e19).  inv_scale == NULL is
 * dgs_step_adam / dgs_step_sgd bit for bit. */
int dgs_step_adam_scaled(void* param, void* exp_avg, void* exp_avg_sq, int state_dtype, const void* grad,
                         int grad_dtype, int64_t n, const dgs_adam_range* ranges, int n_ranges,
                         const float* total_norm, float max_norm, const float* inv_scale,
                         dgs_stream_t stream);
int dgs_step_sgd_scaled(void* param, void* momentum_buf, int state_dtype, const void* grad, int grad_dtype,
                        int64_t n, const dgs_sgd_range* ranges, int n_ranges, const float* total_norm,
                        float max_norm, const float* inv_scale, dgs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * 3b. single-launch step for one GPU: flatten + scale + (global L2 norm) + clip + optimizer step
 *
 * Replaces, on one slot, the same reference lines as sections 1-3 together
 * (H/pytorch/_pytorch_context.py:818-825 _average_gradients, :902 clip_grads == clip_grad_norm_,
 * :914-924 optimizer.step() + zero_grad()) in ONE cooperative kernel that reads the gradient tensors
 * through the segment table instead of a wire buffer: the averaged gradient
 * cast((acc + g) * scale) is formed on the fly (acc == NULL: g * scale), so it is never written and
 * re-read (fp32 AdamW + clip: 32 B/element instead of 40; no clip: 28 instead of 36), and a small
 * model pays one launch instead of four.  seg_range[i] names the row of `ranges` that holds tensor
 * i's hyper-parameters (ranges[].begin/end are ignored here); every tensor of the optimizer must be
 * listed (elements that no segment covers are left untouched).  clip != 0: the global norm of the
 * averaged gradients is computed first (one grid-wide barrier), written to out_norm[0], and every
 * gradient is multiplied by min(max_norm / (norm + 1e-6), 1); `partials` is scratch for
 * partials_capacity floats (>= 1; more allows a larger grid, 148 x 8 is never exceeded).
 * At most DGS_DIRECT_MAX_SEGS tensors per call; gradient dtype must equal the state dtype.
 * ---------------------------------------------------------------------------------------------- */
#define DGS_DIRECT_MAX_SEGS 512

int dgs_direct_step_adam(const dgs_seg* segs, const int32_t* seg_range, int n_segs, int grad_dtype,
                         const float* acc, float scale, void* param, void* exp_avg, void* exp_avg_sq,
                         int state_dtype, const dgs_adam_range* ranges, int n_ranges, int clip,
                         float max_norm, float* partials, int partials_capacity, float* out_norm,
                         dgs_stream_t stream);
/* momentum_buf may be NULL when every range has momentum == 0 */
int dgs_direct_step_sgd(const dgs_seg* segs, const int32_t* seg_range, int n_segs, int grad_dtype,
                        const float* acc, float scale, void* param, void* momentum_buf, int state_dtype,
                        const dgs_sgd_range* ranges, int n_ranges, int clip, float max_norm,
                        float* partials, int partials_capacity, float* out_norm, dgs_stream_t stream);

/* Device-side timing of the optimizer kernels (grad-sync us/step without host noise): while enabled, every
 * dgs_step_adam / dgs_step_sgd launch is bracketed by CUDA events recorded inside the call, on the stream it
 * is launched on.  dgs_timer_read synchronises with the last recorded launch, writes the elapsed microseconds
 * of up to `capacity` launches since the previous read (oldest first) and clears the log (at most 1024 are kept).
 * No reference counterpart: the reference times regions with host clocks
 * (H/profiler.py:487-506 record_timing, optionally after torch.cuda.synchronize). */
int dgs_timer_enable(int enabled);
int dgs_timer_read(float* us_out, int capacity, int* n_out);

/* ------------------------------------------------------------------------------------------------
 * 4. metric reducer
 *
 * Replaces: the per-batch `metric.cpu().detach().numpy()` at H/pytorch/_pytorch_trial.py:920-927,
 * _average_training_metrics (H/pytorch/_metric_utils.py:29-60), the batch mean of
 * util.make_metrics (H/util.py:171-196) and _simple_reduce_metrics (H/pytorch/_reducer.py:28-46).
 * ---------------------------------------------------------------------------------------------- */
typedef struct dgs_scalar_ref {
  const void* ptr; /* device scalar, or NULL for "this batch has no value" (Python None) */
  int32_t dtype;
  int32_t _pad;
} dgs_scalar_ref;

#define DGS_MAX_METRICS 64

/* row[j] = (double)*refs[j].ptr, valid[j] = (refs[j].ptr != NULL); no host sync. refs is a HOST array. */
int dgs_metric_pack(const dgs_scalar_ref* refs, int m, double* row, uint8_t* valid,
                    dgs_stream_t stream);

/* gathered [N][B][M] doubles with validity bytes [N][B][M], num_batches [N] (device int64).
 * Ranks with num_batches == 0 are ignored.  per_batch[b][j] = mean over ranks of the valid values
 * (per_batch_valid 0 if none); avg[j] = mean over b of the valid per_batch values; count[j] = how
 * many batches contributed (int64, exact). */
int dgs_metric_average(const double* gathered, const uint8_t* valid, const int64_t* num_batches,
                       int n_ranks, int b, int m, double* per_batch, uint8_t* per_batch_valid,
                       double* avg, int64_t* count, dgs_stream_t stream);

/* Same reduction over PACKED per-rank records, so that one allgather carries everything a rank reports:
 * record r starts at records + r * record_bytes and holds [ b*m doubles | 1 int64 num_batches | b*m validity
 * bytes ] (record_bytes a multiple of 8, >= 8*b*m + 8 + b*m). */
int dgs_metric_average_packed(const void* records, int64_t record_bytes, int n_ranks, int b, int m,
                              double* per_batch, uint8_t* per_batch_valid, double* avg, int64_t* count,
                              dgs_stream_t stream);

/* Reducer enum values of H/pytorch/_reducer.py:9-25 */
#define DGS_REDUCE_AVG 1
#define DGS_REDUCE_SUM 2
#define DGS_REDUCE_MAX 3
#define DGS_REDUCE_MIN 4

/* values [n][m] (device doubles), weights [n] device int64 or NULL (AVG only: np.average weights).
 * out[j] = reduce over n of values[.][j]. */
int dgs_metric_reduce(const double* values, const int64_t* weights, int n, int m, int op,
                      double* out, dgs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * 5. communicator: NCCL over NVLink on a side stream
 *
 * Replaces: dist.init_process_group (H/pytorch/_pytorch_trial.py:296-300), DDP's bucket allreduce
 * and constructor / per-forward broadcasts (H/pytorch/_pytorch_context.py:296), hvd.broadcast_*
 * (H/pytorch/_pytorch_trial.py:617-620) and the tensor payload of DistributedContext.gather /
 * allgather (H/core/_distributed.py:287-356).
 *
 * libnccl.so.2 is resolved at run time with dlopen (the copy torch already loaded, if any).
 * The 128-byte unique id travels between ranks through the caller's control plane.
 * Collectives run on a high-priority stream owned by the communicator: each *_async call makes
 * that stream wait for everything enqueued so far on `compute`; dgs_comm_wait makes `compute`
 * wait for everything issued so far on the side stream.  Neither blocks the host.
 * One owner thread per communicator.
 * ---------------------------------------------------------------------------------------------- */
typedef struct dgs_comm dgs_comm;

int dgs_comm_unique_id(void* out128);
int dgs_comm_init(const void* id128, int rank, int nranks, int device, dgs_comm** out);
int dgs_comm_destroy(dgs_comm* comm);
int dgs_comm_rank(const dgs_comm* comm, int* rank, int* nranks);
/* in-place sum over ranks of buf[0..count) */
int dgs_comm_allreduce_async(dgs_comm* comm, void* buf, int64_t count, int dtype,
                             dgs_stream_t compute);
int dgs_comm_broadcast_async(dgs_comm* comm, void* buf, int64_t bytes, int root,
                             dgs_stream_t compute);
/* In-place reduce-scatter (sum): buf holds nranks * count_per_rank elements; afterwards this rank's slice
 * buf[rank * count_per_rank ...] holds the sum over ranks of that slice (the other slices are undefined).
 * First half of an allreduce: with dgs_comm_allgather_async (send = recv + rank * bytes_per_rank, in place) on the
 * updated PARAMETERS it lets every rank run the optimizer step on 1/nranks of the elements only. */
int dgs_comm_reduce_scatter_async(dgs_comm* comm, void* buf, int64_t count_per_rank, int dtype,
                                  dgs_stream_t compute);
/* recv holds nranks * bytes_per_rank bytes, ordered by rank */
int dgs_comm_allgather_async(dgs_comm* comm, const void* send, void* recv, int64_t bytes_per_rank,
                             dgs_stream_t compute);
/* Pass 1 of the L2 norm (dgs_l2norm_partial) over x[0..n), enqueued on the communicator's side stream
 * behind the collectives issued so far -- typically right behind the allreduce of the same bucket, while
 * the bucket is still in L2 and the compute stream is busy with backward.  Takes the norm's read of the
 * gradients (H/pytorch/_pytorch_context.py:902) off the critical tail of step_optimizer. */
int dgs_comm_l2norm_partial_async(dgs_comm* comm, const void* x, int dtype, int64_t n, float* partials);
int dgs_comm_wait(dgs_comm* comm, dgs_stream_t compute);
int dgs_nccl_version(int* version);

/* ------------------------------------------------------------------------------------------------
 * 6. peer-memory exchange: the bucket allreduce as this library's own kernels over NVLink / NVSwitch
 *
 * Replaces: DistributedDataParallel's bucket allreduce on the boundary backward pass
 * (H/pytorch/_pytorch_context.py:295-296 wraps the model, :813-816 runs the backward whose hooks
 * reduce the buckets) -- the same call sites as dgs_comm_allreduce_async, for up to DGS_PEER_MAX
 * ranks whose GPUs can map each other's memory (one NVSwitch domain).  Instead of NCCL:
 *   dgs_flatten_push        flatten + scale + cast + reduce-scatter transfer in one kernel: each value
 *                           is stored straight into the wire buffer of the rank that owns its slice
 *   dgs_peer_barrier        all ranks' stores have landed (flags in peer-mapped sync blocks)
 *   dgs_peer_reduce         owner sums its N rows in rank order, emits sum-of-squares partials
 *   dgs_peer_norm_exchange  8-byte all-to-all of the ranks' sums of squares -> global L2 norm
 *   dgs_step_*_push         clip + step on the owner's slice; new parameters stored to every rank
 * Bucket layout: a bucket is nranks * span contiguous elements of the flat buffer, rank j owns
 * elements [begin + j*span, begin + (j+1)*span).  In rank j's OWN wire buffer, row r of the bucket
 * (elements [begin + r*span, ...)) receives rank r's contribution to slice j; after dgs_peer_reduce
 * row j holds the sum, i.e. the reduced gradient sits where the flat layout expects it.
 * All remote accesses are ordinary global stores (and flag loads) on peer-mapped pointers.
 *
 * Memory is shared with CUDA IPC: dgs_peer_export gives the 64-byte handle of the allocation that
 * contains `ptr` and ptr's offset in it; the handles travel through the caller's control plane;
 * dgs_peer_open maps a peer's allocation (once per process, cached) and returns the peer's pointer.
 * The library still allocates no tensor memory.  Sync block: DGS_PEER_SYNC_BYTES of zero-initialised
 * device memory per rank, exported and opened like the buffers.  Flags carry an epoch that the
 * caller increases by one per use of a slot on every rank in lockstep.  A wait that exceeds
 * timeout_s (<= 0: 120 s) records a code readable with dgs_peer_error, stops waiting, and makes every
 * later wait return at once: the stream drains and the host can raise instead of hanging.
 * ---------------------------------------------------------------------------------------------- */
#define DGS_PEER_MAX 8
#define DGS_PEER_SLOTS 4
#define DGS_PEER_HANDLE_BYTES 64
#define DGS_PEER_SYNC_BYTES 512 /* DGS_PEER_SLOTS*DGS_PEER_MAX uint32 flags + 2*DGS_PEER_MAX doubles */

int dgs_peer_export(const void* ptr, void* handle64_out, int64_t* offset_out);
int dgs_peer_open(const void* handle64, int64_t offset, void** out_ptr);
/* one dgs_peer_close per dgs_peer_open of the same handle; the mapping goes when the last one is closed.  The caller
 * makes sure (control plane) that no rank still uses the mapping and that its own kernels on it have finished. */
int dgs_peer_close(const void* handle64);
int dgs_peer_close_all(void);
int dgs_peer_error(int* code_out);
int dgs_peer_error_clear(void);
/* sync_blocks[r] = rank r's sync block as mapped in this process ([rank] is the local one) */
int dgs_peer_barrier(void* const* sync_blocks, int nranks, int rank, int slot, uint32_t epoch,
                     double timeout_s, dgs_stream_t stream);
/* Same modes and arguments as dgs_flatten_accumulate with DGS_WRITE_WIRE set, except that the wire
 * value of flat element f is stored to wire_peers[owner(f)] + f + (rank - owner(f)) * span.  All
 * segments must lie inside [bucket_begin, bucket_begin + nranks*span). */
int dgs_flatten_push(const dgs_seg* segs, int n_segs, int grad_dtype, float* acc, void* const* wire_peers,
                     int wire_dtype, int mode, float scale, int64_t bucket_begin, int64_t span,
                     int nranks, int rank, dgs_stream_t stream);
/* bucket = this rank's wire + bucket_begin.  partials receives dgs_l2norm_grid(span) floats. */
int dgs_peer_reduce(void* bucket, int dtype, int64_t span, int nranks, int rank, float* partials,
                    dgs_stream_t stream);
/* dgs_peer_reduce over every bucket of one flat buffer in ONE launch (a model has a handful of buckets; per-bucket
 * launches are what bounds small models): bucket b is wire[bucket_begin[b] .. + nranks*bucket_span[b]).  Writes one
 * partial per CTA to partials[0..grid) and zero to the rest of partials[0..partials_capacity). */
#define DGS_PEER_MAX_BUCKETS 128
int dgs_peer_reduce_multi(void* wire, int dtype, const int64_t* bucket_begin, const int64_t* bucket_span,
                          int n_buckets, int nranks, int rank, float* partials, int partials_capacity,
                          dgs_stream_t stream);
/* All-gather by stores: for every piece i, bytes [piece_off[i], +piece_bytes[i]) of bases[rank] are copied to the same
 * offsets of bases[r] for every other rank r (offsets and sizes multiples of 16).  The alternative to dgs_step_*_push
 * for the parameter exchange: dgs_step_* on the local slice, then this. */
int dgs_peer_scatter(void* const* bases, const int64_t* piece_off, const int64_t* piece_bytes, int n_pieces,
                     int nranks, int rank, dgs_stream_t stream);
/* out_norm[0] = sqrt(sum over ranks (in rank order) of sum(partials)); implies a barrier */
int dgs_peer_norm_exchange(void* const* sync_blocks, int nranks, int rank, int slot, uint32_t epoch,
                           const float* partials, int n_partials, float* out_norm, double timeout_s,
                           dgs_stream_t stream);
/* dgs_step_adam / dgs_step_sgd on this rank's slice (param, state, grad point at the slice) with the
 * updated parameters stored to param_peers[0..nranks) (every rank's pointer to the same slice;
 * param_peers[rank] == param). */
int dgs_step_adam_push(void* param, void* exp_avg, void* exp_avg_sq, int state_dtype, const void* grad,
                       int grad_dtype, int64_t n, const dgs_adam_range* ranges, int n_ranges,
                       const float* total_norm, float max_norm, void* const* param_peers, int nranks,
                       int rank, dgs_stream_t stream);
int dgs_step_sgd_push(void* param, void* momentum_buf, int state_dtype, const void* grad, int grad_dtype,
                      int64_t n, const dgs_sgd_range* ranges, int n_ranges, const float* total_norm,
                      float max_norm, void* const* param_peers, int nranks, int rank, dgs_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DGS_H_ */
