"""GradSyncEngine: the host side of the B200 grad-sync path, one engine per wrapped optimizer.

What the reference does with torch DDP / Horovod + torch.optim between `loss.backward()` and the end
of `step_optimizer()` (reference harness/determined/pytorch/_pytorch_context.py:718-924) happens
here as calls into libdgs.so:

  backward, per bucket, as soon as the bucket's last gradient is produced (autograd hook):
      dgs_flatten_accumulate   grads -> fp32 accumulator (non-boundary micro-batches)
                               grads (+ accumulator) -> wire bucket * 1/(N*A) (boundary micro-batch)
      dgs_comm_allreduce_async wire bucket, on the NCCL side stream, overlapped with the rest of backward
  step_optimizer:
      dgs_comm_wait            compute stream waits for the side stream (no host block)
      dgs_l2norm_partial/final global norm when the user's clip_grads is clip_grad_norm_
      dgs_step_adam / _sgd     clip + optimizer step over the flat buffers

Data layout in HBM, per parameter dtype of the optimizer ("flat set"):
  param   [total]  param dtype   every p.data is a view into it (state_dict / checkpoints unchanged)
  wire    [total]  wire dtype    averaged gradients; param dtype, or fp16 under gradient_compression
  acc     [total]  fp32          only when gradients are aggregated over several backward passes
  state   [total] x {1,2}        exp_avg / exp_avg_sq or momentum_buffer; optimizer.state[p][...] are views
Tensors are laid out in optimizer order (param_groups flattened), each starting on a 128-byte
boundary, so a param group is one contiguous element range and a comm bucket is one contiguous
slice.  Buckets are cut from the END of the buffer (gradients arrive roughly in reverse), 1 MiB
first then 25 MiB, the sizes torch DDP uses.

Two shortcuts replace parts of that sequence where nothing needs the averaged gradient in memory:
  * one GPU, no accumulation: the final pass's gradients stay where autograd left them and step_optimizer runs
    dgs_direct_step_* -- scale, norm, clip and optimizer step in ONE kernel (direct_step / flush_deferred);
  * several GPUs on one NVSwitch box, route "peer" (what route "auto" moves to): dgs_flatten_push stores each value into
    the wire buffer of the rank that owns its slice, the owner reduces (dgs_peer_reduce_multi), the ranks exchange
    their sums of squares (dgs_peer_norm_exchange) and dgs_step_*_push steps the own slices and stores the new
    parameters to every rank -- no NCCL call in a steady-state step (_peer.py, csrc/dgs_peer.cu).

There is no CPU or torch fallback for these steps.  Optimizers other than SGD / Adam / AdamW keep
their own torch `step()` on gradient views of the wire buffer (still flattened, all-reduced and
clipped by the kernels).
"""
import ctypes
import os as _os
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import torch

from determined_clone_b200 import _lib, errors
from determined_clone_b200.pytorch import _peer
from determined_clone_b200.pytorch._comm import Comm

ALIGN_ELEMS = 64  # 128 bytes of bf16, 256 bytes of fp32: every tensor starts vector-aligned
FIRST_BUCKET_BYTES = 1 << 20
# Bucket cap: NCCL's allreduce over NVSwitch needs large messages (measured at 8 x B200, gpurun_out/
# nccl_probe_n8.jsonl: bus bandwidth 331 GB/s at 25 MiB, 418 at 64, 590 at 128, 728 at 1 GiB), small models
# want several buckets to overlap with backward.  Default: 1/8 of the gradient bytes clamped to
# [25 MiB (torch DDP's size), 128 MiB]; DGS_BUCKET_MB fixes it for tuning.
_BUCKET_MB_ENV = __import__("os").environ.get("DGS_BUCKET_MB")
BUCKET_BYTES = (int(_BUCKET_MB_ENV) << 20) if _BUCKET_MB_ENV else 0  # 0 = adaptive
MIN_BUCKET_BYTES, MAX_BUCKET_BYTES = 25 << 20, 128 << 20
# Experiment switch, off by default: flat sets smaller than this many MiB become ONE bucket --
# one allreduce per step instead of a 1 MiB bucket plus 25 MiB ones -- for launch-bound small models, where each
# NCCL enqueue costs the host more than the lost overlap costs the GPU.
# Peer route, parameter exchange: the step kernel stores the new parameters to every rank itself ("fused"), or steps the
# local slice and a copy kernel spreads it ("split").  DGS_PUSH_MODE picks; see DESIGN section 7 for the measurement.
_PUSH_SPLIT = _os.environ.get("DGS_PUSH_MODE", "fused") == "split"
_SLICE_LAUNCH = _os.environ.get("DGS_SLICE_LAUNCH", "single")
# Peer route: flat sets whose largest bucket reaches this size reduce every bucket on a side stream as soon as ITS
# pushes have landed everywhere (barrier + owner reduction per bucket, hidden behind the pushes of the buckets that
# follow); smaller sets keep ONE barrier and ONE reduction launch at step time (fewer launches matter more there).
# DGS_PEER_OVERLAP_MB=0 switches the per-bucket mode off.
_PEER_OVERLAP_MIN = int(_os.environ.get("DGS_PEER_OVERLAP_MB", "32")) << 20
ONE_BUCKET_BELOW_BYTES = int(__import__("os").environ.get("DGS_ONE_BUCKET_BELOW_MB", "0")) << 20


def _check_device(device: torch.device) -> None:
    if device.type != "cuda":
        raise _lib.DgsError("GradSyncEngine needs a CUDA device: the grad-sync path has no CPU implementation")


def _load_host_ext() -> Any:
    """The native host loop (csrc/dgs_host.cpp: gradient pick-up and segment tables in C++), used whenever it has been
    built (__graft_entry__.build() / _build.build_host()); DGS_HOST_EXT=0 keeps the Python loops, =1 insists on it.
    Measured on B200 (profiles/r2_host_ext_ab.txt): 16 MiB / 64 tensors, fp32 AdamW + clip, one GPU: 133 -> 93 us per
    optimizer step at A=1, 244 -> 156 us at A=4."""
    want = __import__("os").environ.get("DGS_HOST_EXT", "auto")
    if want == "0":
        return None
    import importlib.util

    from determined_clone_b200 import _build

    path = _build.host_ext_path()
    if not path.exists() and want != "1":
        return None  # not built (python -m determined_clone_b200._build builds libdgs only): the Python loops do the same work
    if not path.exists():
        raise _lib.DgsError(f"DGS_HOST_EXT=1 but {path} is not built (python -c 'from determined_clone_b200 import _build; _build.build_host()')")
    spec = importlib.util.spec_from_file_location("_dgs_host", str(path))
    mod = importlib.util.module_from_spec(spec)  # type: ignore
    spec.loader.exec_module(mod)  # type: ignore
    return mod


_host = _load_host_ext()


def _align(n: int) -> int:
    return (n + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS


try:
    _raw_stream = torch._C._cuda_getCurrentRawStream  # ~0.2 us; torch.cuda.current_stream() costs ~5 us
except AttributeError:  # pragma: no cover
    _raw_stream = None


def plan_layout(numels: Sequence[int]) -> Tuple[List[int], int]:
    """Element offset of every tensor in a flat buffer (each on an ALIGN_ELEMS boundary) and the buffer length."""
    offs, total = [], 0
    for n in numels:
        offs.append(total)
        total = _align(total + n)
    return offs, total


def plan_buckets(offs: Sequence[int], numels: Sequence[int], total: int, itemsize: int, first_bytes: int,
                 cap_bytes: int) -> List[Tuple[int, int, List[int]]]:
    """Cut the flat buffer into communication buckets (begin, end, tensor indices), from the END: the first
    bucket closes once it holds first_bytes, the following ones at cap_bytes; a bucket is a contiguous element
    range made of whole tensors, and together the buckets tile [0, total)."""
    buckets: List[Tuple[int, int, List[int]]] = []
    limit, cur, cur_bytes, end = first_bytes, [], 0, total
    for i in range(len(offs) - 1, -1, -1):
        cur.append(i)
        cur_bytes += numels[i] * itemsize
        if cur_bytes >= limit:
            buckets.append((offs[i], end, cur))
            end, cur, cur_bytes, limit = offs[i], [], 0, cap_bytes
    if cur:
        buckets.append((0, end, cur))
    return buckets


def adaptive_bucket_cap(total_bytes: int) -> int:
    return min(max(total_bytes // 8, MIN_BUCKET_BYTES), MAX_BUCKET_BYTES)


def _stream() -> ctypes.c_void_p:
    """cudaStream_t of torch's current stream on the current device."""
    if _raw_stream is not None:
        return ctypes.c_void_p(_raw_stream(torch.cuda.current_device()))
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t: Optional[torch.Tensor], byte_off: int = 0) -> ctypes.c_void_p:
    return ctypes.c_void_p(None if t is None else t.data_ptr() + byte_off)


class ClipRequest:
    """What the user's clip_grads asked for, when it is torch.nn.utils.clip_grad_norm_(all params, max_norm)."""

    def __init__(self) -> None:
        self.max_norm: Optional[float] = None


class _Tensor:
    __slots__ = ("param", "group", "off", "numel", "bucket", "index")

    def __init__(self, param: torch.nn.Parameter, group: int, off: int, index: int) -> None:
        self.param, self.group, self.off, self.numel, self.index = param, group, off, param.numel(), index
        self.bucket = -1


class _Bucket:
    __slots__ = ("begin", "end", "tensors", "flushed", "reduced", "norm_off", "norm_n", "norm_done", "segs", "params", "every_tensor")

    def __init__(self, begin: int, end: int, tensors: List[int]) -> None:
        self.begin, self.end, self.tensors = begin, end, tensors
        self.norm_off = self.norm_n = 0  # this bucket's slots in the engine's norm partials
        self.norm_done = False  # pass 1 of the norm already ran on this bucket in this window
        self.segs: Any = None  # prebuilt dgs_seg[len(tensors)] (dst_off, numel fixed; src set per flush)
        self.params: List[Any] = []  # the bucket's parameters, in table order
        self.every_tensor: Dict[int, Any] = {}  # {tensor index: None}: "all of them" for _flush_bucket
        self.flushed = False  # wire produced for this window
        self.reduced = False  # allreduce issued for this window


class _FlatSet:
    """All parameters of one dtype of one optimizer."""

    def __init__(self, dtype: torch.dtype, device: torch.device) -> None:
        self.dtype, self.device = dtype, device
        self.wire_dtype = dtype
        self.tensors: List[_Tensor] = []
        self.buckets: List[_Bucket] = []  # in launch order: buckets[0] covers the END of the buffer
        self.total = 0
        self.param: torch.Tensor = None  # type: ignore
        self.wire: torch.Tensor = None  # type: ignore
        self.acc: Optional[torch.Tensor] = None
        self.state: Dict[str, torch.Tensor] = {}
        self.state_views: Dict[str, List[torch.Tensor]] = {}
        self.grad_views: List[torch.Tensor] = []
        self.in_window: List[bool] = []  # tensor has contributed to the accumulator in this window
        self.in_wire: List[bool] = []  # tensor's wire slice is final for this window
        self.dtype_code = self.wire_code = 0
        self.acc_ptr: Any = None
        self.wire_ptr: Any = None
        # single GPU: the final pass's gradients, kept (not yet flattened) until step_optimizer shows whether the
        # single-launch step can take them (dgs_direct_step_*): (gradient tensors, accumulator already holds earlier passes)
        self.deferred: Optional[Tuple[List[torch.Tensor], bool]] = None
        self.seg_range_for: Any = None  # which run structure seg_range currently describes
        self.seg_range: Any = None  # (int32)[len(tensors)]: hyper-parameter row of every tensor, for the direct step
        self.peer_overlap = False  # peer route: barrier + owner reduction per bucket on the side stream (large buckets)
        self.scatter_table: Any = None  # peer route, split parameter exchange: (offsets, sizes, n) of dgs_peer_scatter
        self.reduce_table: Any = None  # peer route: (begins, spans, first norm slot, slots) of dgs_peer_reduce_multi
        self.peer_wire: Any = None  # (void*)[world]: every rank's pointer to its wire buffer of this set (peer route)
        self.segs_all: Any = None  # prebuilt dgs_seg[len(tensors)] for whole-set launches (single GPU)
        self.params: List[Any] = []  # every parameter of the set, in table order
        self.group_runs: Optional[List[Tuple[int, int, int, List[_Tensor]]]] = None
        self.steady = False  # last step covered every parameter and left every state tensor as our view
        self.uniform_step: Optional[int] = None  # Adam: the common step count while all parameters move together


class GradSyncEngine:
    def __init__(
        self,
        optimizer: torch.optim.Optimizer,
        device: torch.device,
        comm: Optional[Comm],
        aggregation_frequency: int,
        average_aggregated_gradients: bool,
        backward_passes_per_step: int = 1,
        fp16_compression: bool = False,
        bucket_bytes: int = BUCKET_BYTES,
        first_bucket_bytes: int = FIRST_BUCKET_BYTES,
        sharded_step: Optional[bool] = None,
        route: Optional[str] = None,
    ) -> None:
        _lib.load()  # fail now, loudly, if libdgs.so is missing
        _check_device(device)
        self.optimizer = optimizer
        self.device = device
        self.comm = comm
        self.world = comm.size if comm is not None else 1
        self.agg = aggregation_frequency
        self.average = average_aggregated_gradients
        self.bpps = backward_passes_per_step
        self.fp16_compression = fp16_compression
        self.bucket_bytes, self.first_bucket_bytes = bucket_bytes, first_bucket_bytes
        self.scale = 1.0 / (self.world * (self.agg if self.average else 1))
        # extra factor folded into the flatten kernel's scale: 1/S while an auto-AMP GradScaler's loss scale S is
        # in the gradients (PyTorchTrialContext.backward sets it per window), else exactly 1.0
        self.unscale = 1.0
        self.stepped_in_batch: Optional[int] = None  # batch index of the last optimizer step (PyTorchTrialContext.backward)
        self.needs_acc = self.agg * self.bpps > 1
        self.sets: List[_FlatSet] = []
        self._where: Dict[int, Tuple[int, int]] = {}  # id(param) -> (set index, tensor index)
        self._n_group_params = 0  # parameters in the optimizer's groups when it was wrapped
        self._frozen_at_wrap: List[torch.nn.Parameter] = []  # of those, the ones left out (requires_grad=False)
        self._hooks: List[Any] = []  # autograd hooks (order learning, then one sentinel per bucket)
        self._state_hook: Any = None
        self._sentinels: Optional[Dict[int, Tuple[int, int]]] = None  # id(param) -> (set, bucket); None = not learnt
        self._order: List[Tuple[int, int]] = []
        self._learn_tries = 0
        self._final_pass = False
        self._in_pass = False
        self._norm = torch.zeros(1, dtype=torch.float32, device=device)
        self._partials: torch.Tensor = None  # type: ignore  # scratch of the L2-norm kernels, sized in _build
        self._steps_host: Optional[torch.Tensor] = None  # Adam's per-parameter `step`, one host tensor
        self._step_views: List[torch.Tensor] = []
        self._fused_kind = self._classify(optimizer)
        # Exchange route of the optimizer step on several GPUs:
        #   allreduce  NCCL allreduce of each bucket, every rank steps every element
        #   sharded    NCCL reduce-scatter, every rank clips and steps only its 1/N slice of each bucket, NCCL
        #              all-gather of the updated parameters (round 1's opt-in sharded step)
        #   peer       the same slicing, but the exchange is this library's own kernels on peer-mapped buffers over
        #              NVLink: the flatten kernel stores each value into the owner's wire buffer, the owner reduces,
        #              the step kernel stores the new parameters to every rank (csrc/dgs_peer.cu, _peer.py)
        #   auto       (default) start on allreduce; once two optimizer steps in a row used the path in the way the
        #              sliced routes can honour (no GradScaler, clip_grads absent or provably nothing but
        #              clip_grad_norm_, auto_zero_grads) move to peer for good -- if peer set-up succeeded on every
        #              rank.  A later step that does not fit moves back to allreduce (demote()).
        # With sliced routes the optimizer state of other ranks' slices is fetched when optimizer.state_dict() is called.
        if route is None:
            route = "sharded" if sharded_step else _os.environ.get("DGS_ROUTE") or ("sharded" if _os.environ.get("DGS_SHARDED_STEP", "") not in ("", "0") else "auto")
        if route not in ("auto", "allreduce", "sharded", "peer"):
            raise errors.InvalidExperimentException(f"optimizations.grad_sync_route must be auto, allreduce, sharded or peer; got {route!r}")
        sliceable = bool(comm is not None and self._fused_kind is not None and not fp16_compression and ALIGN_ELEMS % (8 * self.world) == 0)
        if route in ("sharded", "peer") and comm is not None and not sliceable:
            route = "allreduce"  # non-fused optimizer, fp16 wire compression or a world size the 128-byte alignment does not divide
        self.route = route if comm is not None else "allreduce"
        self.sliceable = sliceable
        self.sharded = self.route in ("sharded", "peer")  # this rank holds (and steps) only its slice of every bucket
        self.peer = self.route == "peer"                   # ... exchanged by the peer kernels instead of NCCL
        # One GPU: flatten + norm + clip + step as ONE kernel reading the gradient tensors in place (csrc/dgs_direct.cu)
        # whenever the optimizer step is ours and nothing needs the averaged gradients in memory; DGS_DIRECT=0 keeps
        # the three-kernel path (flatten -> wire, norm, step).
        self.direct = comm is None and self._fused_kind is not None and _os.environ.get("DGS_DIRECT", "1") != "0"
        self._direct_clip: Optional[Tuple[bool, float]] = None  # set for the duration of direct_step()
        self._all_present: Optional[List[List[bool]]] = None
        self._peer_stream: Any = None  # side stream of the per-bucket owner reductions (peer route, large buckets)
        self._peer_side_pending = False
        self.peers: Optional[_peer.PeerGroup] = None
        self.peer_unavailable: Optional[str] = None  # why peer set-up failed, if it did
        self._conforming_steps = 0
        self.last_total_norm: Optional[torch.Tensor] = None
        self._inv_scale: Optional[torch.Tensor] = None
        # Several GPUs: norm pass 1 is run per bucket right behind the bucket's allreduce (aux stream, bucket still
        # in L2) once a clip_grad_norm_ request has been seen; switched off again if nobody asks for the norm.
        # One GPU: a single pass over the whole buffer at step time (fewer launches, nothing to hide behind).
        self._eager_norm = False
        self._eager_unused = 0
        self._norm_slots_valid = True
        self._norm_requested = False
        self._wire_versions: List[int] = []
        self._launch_order: List[Tuple[int, int]] = []
        self._pending_order: Optional[List[Tuple[int, int]]] = None
        self._next_launch = 0
        self._build()

    # ---- construction ---------------------------------------------------------------------------
    @staticmethod
    def _classify(opt: torch.optim.Optimizer) -> Optional[str]:
        """Which fused kernel can stand in for opt.step(); None keeps torch's own step."""
        kind = {torch.optim.SGD: "sgd", torch.optim.Adam: "adam", torch.optim.AdamW: "adamw"}.get(type(opt))
        if kind is None:
            return None
        for g in opt.param_groups:
            if g.get("differentiable") or g.get("capturable") or g.get("amsgrad"):
                return None
            if isinstance(g.get("lr"), torch.Tensor):
                return None
            if kind != "sgd" and any(isinstance(b, torch.Tensor) for b in g["betas"]):
                return None
        return kind

    def _build(self) -> None:
        by_dtype: Dict[torch.dtype, _FlatSet] = {}
        for gi, group in enumerate(self.optimizer.param_groups):
            for p in group["params"]:
                self._n_group_params += 1
                if not p.requires_grad:
                    self._frozen_at_wrap.append(p)
                    continue
                if p.device != self.device:
                    raise errors.InvalidExperimentException(
                        f"optimizer parameter on {p.device}, expected {self.device}: wrap_model() must be "
                        "called before the optimizer is created and wrapped"
                    )
                if p.dtype not in (torch.float32, torch.bfloat16, torch.float16):
                    raise _lib.DgsError(f"parameter dtype {p.dtype} is not supported by the grad-sync kernels")
                fs = by_dtype.get(p.dtype)
                if fs is None:
                    fs = by_dtype[p.dtype] = _FlatSet(p.dtype, self.device)
                    self.sets.append(fs)
                if id(p) in self._where:
                    continue  # torch rejects duplicates across groups already
                t = _Tensor(p, gi, fs.total, len(fs.tensors))
                fs.total = _align(fs.total + t.numel)
                self._where[id(p)] = (self.sets.index(fs), t.index)
                fs.tensors.append(t)
        for fs in self.sets:
            if self.fp16_compression and fs.dtype == torch.float32:
                fs.wire_dtype = torch.float16  # hvd.Compression.fp16 (_pytorch_context.py:357-366)
            fs.param = torch.zeros(fs.total, dtype=fs.dtype, device=self.device)
            fs.wire = torch.zeros(fs.total, dtype=fs.wire_dtype, device=self.device)
            if self.needs_acc:
                fs.acc = torch.zeros(fs.total, dtype=torch.float32, device=self.device)
            for t in fs.tensors:
                view = fs.param[t.off:t.off + t.numel].view(t.param.shape)
                view.copy_(t.param.data)
                t.param.data = view
                fs.grad_views.append(fs.wire[t.off:t.off + t.numel].view(t.param.shape))
            fs.in_window = [False] * len(fs.tensors)
            fs.in_wire = [False] * len(fs.tensors)
            self._cut_buckets(fs)
            fs.dtype_code, fs.wire_code = _lib.dtype_code(fs.dtype), _lib.dtype_code(fs.wire_dtype)
            fs.acc_ptr, fs.wire_ptr = _ptr(fs.acc), _ptr(fs.wire)
            fs.segs_all = (_lib.Seg * len(fs.tensors))()
            fs.params = [t.param for t in fs.tensors]
            for k, t in enumerate(fs.tensors):
                fs.segs_all[k].dst_off, fs.segs_all[k].numel = t.off, t.numel
        # norm scratch: [whole-buffer partials of every set | per-bucket partials of every set].  A bucket's slots are
        # sized for the whole bucket, so that they also hold the (fewer) partials of one slice after a route change;
        # unused slots stay zero (_set_route clears the scratch), the final pass sums them all.
        n_partials = sum(_lib.calls.dgs_l2norm_grid(fs.total) for fs in self.sets)
        self._bucket_partials_base = n_partials
        for fs in self.sets:
            for b in fs.buckets:
                b.norm_off, b.norm_n = n_partials, _lib.calls.dgs_l2norm_grid(b.end - b.begin)
                n_partials += b.norm_n
        self._partials = torch.zeros(max(n_partials, 1), dtype=torch.float32, device=self.device)
        self._launch_order = [(si, bi) for si, fs in enumerate(self.sets) for bi in range(len(fs.buckets))]
        if self.comm is not None:
            # what DDP's constructor does: every replica starts from rank 0's parameters
            for fs in self.sets:
                self.comm.broadcast(fs.param, 0)
            self.comm.wait()
            self._state_hook = self.optimizer.register_state_dict_pre_hook(lambda opt: self.gather_state())
        if self.comm is not None and self.sliceable and self.route in ("auto", "peer"):
            self._setup_peers(required=self.route == "peer")

    def _setup_peers(self, required: bool) -> None:
        """Map every rank's wire and parameter buffers (all-or-none across ranks)."""
        comm = self.comm
        assert comm is not None
        try:
            if comm.allgather_obj is None:
                raise _lib.DgsError("the communicator has no control-plane allgather to carry the IPC handles")
            buffers = {}
            for si, fs in enumerate(self.sets):
                buffers[f"wire{si}"], buffers[f"param{si}"] = fs.wire, fs.param
            timeout = float(_os.environ.get("DGS_PEER_TIMEOUT_S", "0") or 0)
            self.peers = _peer.PeerGroup(comm.rank, comm.size, self.device, comm.allgather_obj, buffers, timeout_s=timeout)
            for si, fs in enumerate(self.sets):
                fs.peer_wire = self.peers.table(f"wire{si}")
                biggest = max((b.end - b.begin) * fs.wire.element_size() for b in fs.buckets)
                # (with one big bucket behind the 1 MiB first one there is nothing to hide a reduction behind)
                fs.peer_overlap = _PEER_OVERLAP_MIN > 0 and biggest >= _PEER_OVERLAP_MIN and len(fs.buckets) >= 3
            if self.device.type == "cuda" and any(fs.peer_overlap for fs in self.sets):
                self._peer_stream = torch.cuda.Stream(device=self.device, priority=-1)
        except _lib.DgsError as e:
            self.peers = None
            if required:
                raise
            self.peer_unavailable = str(e)

    def _set_route(self, sliced: bool, peer: bool) -> None:
        """Switch between the full-gradient route and the sliced ones; only between windows."""
        self.sharded, self.peer = sliced, peer
        self._partials.zero_()
        if peer:
            # On the peer route the rows of a bucket in this rank's wire buffer receive the OTHER ranks' contributions
            # to this rank's slice; nobody ever writes the alignment gaps of those rows, and the owner sums them into
            # the norm.  Coming from the allreduce route the same memory holds last step's reduced gradients of other
            # slices: clear it once (found as a 7e-5 parameter drift against the reference's 2-slot fixture).
            for fs in self.sets:
                fs.wire.zero_()
        self._eager_norm = False
        self.last_total_norm = None

    def note_step(self, conforming: bool) -> None:
        """Called by the context after every optimizer step with whether that step used the path in a way the sliced
        routes can honour.  route == auto: two such steps in a row move the engine to the peer route."""
        if self.route != "auto" or self.peers is None:
            return
        self._conforming_steps = self._conforming_steps + 1 if conforming else 0
        if not self.sharded and self._conforming_steps >= 2:
            self._set_route(True, True)

    def demote(self, present: Optional[List[List[bool]]] = None) -> None:
        """A step the sliced routes cannot honour arrived (route == auto only): bring every rank's optimizer state and,
        if a window is waiting (finish_window already ran), the reduced gradient slices back to all ranks, and
        continue on the allreduce route for good."""
        assert self.comm is not None and self.sharded
        self.gather_state()
        if present is not None:
            for fs in self.sets:
                isz = fs.wire.element_size()
                for b in sorted(fs.buckets, key=lambda x: x.begin):
                    self.comm.allgather_inplace_raw(fs.wire.data_ptr() + b.begin * isz, (b.end - b.begin) // self.world * isz)
            self.comm.wait()
        self.route = "allreduce"
        self._set_route(False, False)

    def _cut_buckets(self, fs: _FlatSet) -> None:
        itemsize = torch.empty(0, dtype=fs.wire_dtype).element_size()
        cap = self.bucket_bytes or adaptive_bucket_cap(fs.total * itemsize)
        first = self.first_bucket_bytes
        if fs.total * itemsize < ONE_BUCKET_BELOW_BYTES:
            first = cap = fs.total * itemsize + 1  # never reached: everything lands in the one trailing bucket
        for begin, end, idxs in plan_buckets([t.off for t in fs.tensors], [t.numel for t in fs.tensors], fs.total, itemsize,
                                             first, cap):
            fs.buckets.append(_Bucket(begin, end, idxs))
        for bi, b in enumerate(fs.buckets):
            b.segs = (_lib.Seg * len(b.tensors))()
            b.params = [fs.tensors[ti].param for ti in b.tensors]
            b.every_tensor = dict.fromkeys(b.tensors)
            for k, ti in enumerate(b.tensors):
                t = fs.tensors[ti]
                t.bucket = bi
                b.segs[k].dst_off, b.segs[k].numel = t.off, t.numel

    def close(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks = []
        if self._state_hook is not None:
            self._state_hook.remove()
            self._state_hook = None
        if self.peers is not None:
            self.peers.close()  # unmap the other ranks' buffers; the engine is not usable for peer steps afterwards
            self.peers = None

    # ---- backward ---------------------------------------------------------------------------------
    # Gradients are left where autograd puts them (p.grad) and picked up per bucket:
    #   * single GPU: after loss.backward() returns (end_pass) -- there is nothing to overlap with, and
    #     no Python runs per parameter during backward;
    #   * several GPUs: a hook on ONE parameter per bucket (the one whose gradient is produced last,
    #     learnt from the first backward pass, as torch DDP rebuilds its buckets after the first
    #     iteration) flushes the bucket and starts its allreduce while backward continues.
    def begin_pass(self, final: bool) -> None:
        """Called right before loss.backward().  `final`: this pass completes the window, so finished
        buckets are scaled, cast and all-reduced as soon as they are complete."""
        for fs in self.sets:
            for b in fs.buckets:
                if b.flushed:
                    raise errors.InvalidExperimentException(
                        "context.backward() was called again after the gradients of this batch were "
                        "already synchronized; pass the number of backward calls per batch as "
                        "backward_passes_per_step to wrap_optimizer()"
                    )
        self._in_pass, self._final_pass = True, final
        if self.comm is not None and self._sentinels is None and not self._hooks:
            self._order = []
            for fs in self.sets:
                for t in fs.tensors:
                    self._hooks.append(t.param.register_post_accumulate_grad_hook(self._learn_order))

    def end_pass(self) -> None:
        """Called right after loss.backward(): flatten whatever the hooks have not taken yet."""
        for si, fs in enumerate(self.sets):
            if self.comm is None and self._flush_whole_set(fs):
                continue
            for bi, b in enumerate(fs.buckets):
                grads = self._collect(fs, b)
                if grads:
                    self._flush_bucket(si, bi, grads, self._final_pass and len(grads) == len(b.tensors))
        if self._sentinels is None and self._hooks:
            self._finish_learning()
        self._in_pass = self._final_pass = False

    def _flush_whole_set(self, fs: _FlatSet) -> bool:
        """Single GPU, the usual case (every parameter has a gradient, all in the same accumulation state):
        one call for the whole flat set instead of one per bucket -- buckets only exist to pipeline
        communication.  Returns False, having touched nothing, when the case does not apply."""
        tensors, in_window = fs.tensors, fs.in_window
        state = in_window[0]
        segs = fs.segs_all
        if _host is not None:
            if any(s != state for s in in_window) or any(b.flushed for b in fs.buckets):
                return False
            grads = _host.take_grads(fs.params, ctypes.addressof(segs), fs.dtype_code)  # checks, table and p.grad = None natively
            if grads is None:
                return False
        else:
            grads = [t.param.grad for t in tensors]
            dtype = fs.dtype
            for g in grads:
                if g is None or g.dtype != dtype or g.is_sparse or not g.is_contiguous():
                    return False
            if any(s != state for s in in_window) or any(b.flushed for b in fs.buckets):
                return False
            for k, g in enumerate(grads):  # `grads` keeps the tensors alive until the launch below
                tensors[k].param.grad = None
                segs[k].src = g.data_ptr()
        to_wire = self._final_pass
        if not to_wire and fs.acc is None:
            fs.acc = torch.zeros(fs.total, dtype=torch.float32, device=self.device)
            fs.acc_ptr = _ptr(fs.acc)
        # (only without an accumulator: with one, reading acc + g twice costs what writing and re-reading the wire does,
        # and the three kernels ran closer to the HBM peak -- measured on GPT-2 medium, bf16, A=4: 1.42 ms against 1.28)
        if to_wire and not state and self.direct and len(tensors) <= _lib.DIRECT_MAX_SEGS and fs.wire_dtype == fs.dtype:
            # leave the gradients where they are (the segment table points at them): step_optimizer decides between the
            # single-launch step and this flatten (flush_deferred)
            fs.deferred = (grads, bool(state))
        else:
            mode = (_lib.READ_ACC if state else 0) | (_lib.WRITE_WIRE if to_wire else 0)
            _lib.calls.dgs_flatten_accumulate(segs, len(tensors), fs.dtype_code, fs.acc_ptr, fs.wire_ptr, fs.wire_code, mode,
                                              self.scale * self.unscale, _stream())
        flags = fs.in_wire if to_wire else fs.in_window
        for i in range(len(tensors)):
            flags[i] = True
        if to_wire:
            for b in fs.buckets:
                b.flushed = True
            self._reduce_ready()
        return True

    def _learn_order(self, p: torch.nn.Parameter) -> None:
        self._order.append(self._where[id(p)])

    def _finish_learning(self) -> None:
        """After a backward pass in which every parameter produced exactly one gradient: keep a hook only
        on the last-arriving parameter of each bucket."""
        n = sum(len(fs.tensors) for fs in self.sets)
        self._learn_tries += 1
        complete = len(self._order) == n and len(set(self._order)) == n
        # The ranks must take the same branch below (one of them issues a broadcast): `complete` is rank-local --
        # with a data-dependent graph one rank can miss a parameter that the others saw -- so agree on it first.
        # Every rank gets here in the same backward pass (hooks are installed by the first begin_pass everywhere).
        flag = torch.tensor([1 if complete else 0], dtype=torch.int64, device=self.device)
        self.comm.allreduce(flag)  # type: ignore
        self.comm.wait()  # type: ignore
        complete = int(flag.item()) == self.world
        if not complete and self._learn_tries < 3:
            self._order = []
            return
        for h in self._hooks:
            h.remove()
        self._hooks = []
        self._sentinels = {}
        if not complete:
            return  # irregular graphs: no overlap, buckets are flushed at the end of backward
        last: Dict[Tuple[int, int], int] = {}
        for pos, (si, ti) in enumerate(self._order):
            last[(si, self.sets[si].tensors[ti].bucket)] = pos
        # buckets are all-reduced in the order in which they complete (a param group that collects every
        # layer's 1-D parameters completes last wherever it sits in the buffer); rank 0's order is adopted
        # by everyone so that the NCCL sequences cannot differ
        order = sorted(last, key=lambda k: last[k])
        # complete on every rank => every bucket appears in `order` on every rank: the broadcast below is collective
        assert len(order) == len(self._launch_order)
        flat = torch.tensor([x for sb in order for x in sb], dtype=torch.int64, device=self.device)
        self.comm.broadcast(flat, 0)  # type: ignore
        self.comm.wait()  # type: ignore
        got = flat.tolist()
        # adopted at the next window boundary: this window may already be part-way through the old order
        self._pending_order = [(got[i], got[i + 1]) for i in range(0, len(got), 2)]
        for (si, bi), pos in last.items():
            _, ti = self._order[pos]
            param = self.sets[si].tensors[ti].param
            self._sentinels[id(param)] = (si, bi)
            self._hooks.append(param.register_post_accumulate_grad_hook(self._on_sentinel))

    def _on_sentinel(self, p: torch.nn.Parameter) -> None:
        if not self._final_pass:
            return  # nothing to communicate on this pass; end_pass() flattens everything at once
        si, bi = self._sentinels[id(p)]  # type: ignore
        fs = self.sets[si]
        b = fs.buckets[bi]
        if _host is not None and not b.flushed:
            first_state = fs.in_window[b.tensors[0]]
            if all(fs.in_window[ti] == first_state for ti in b.tensors):
                taken = _host.take_grads(b.params, ctypes.addressof(b.segs), fs.dtype_code)
                if taken is not None:  # else: incomplete bucket or gradients that need converting -> the general path below
                    self._flush_bucket(si, bi, b.every_tensor, True, table_filled=True)
                    return
        for ti in b.tensors:
            if fs.tensors[ti].param.grad is None:
                return  # arrival order changed: leave the bucket to end_pass()
        self._flush_bucket(si, bi, self._collect(fs, b), True)

    def _collect(self, fs: _FlatSet, b: _Bucket) -> Dict[int, torch.Tensor]:
        """Take the gradients autograd has produced for this bucket (p.grad -> None)."""
        out: Dict[int, torch.Tensor] = {}
        tensors = fs.tensors
        for ti in b.tensors:
            p = tensors[ti].param
            g = p.grad
            if g is None:
                continue
            p.grad = None  # the kernel consumes it; autograd must not accumulate into it again
            if g.is_sparse:
                raise _lib.DgsError("sparse gradients are not supported by the grad-sync kernels")
            if g.dtype != fs.dtype or not g.is_contiguous():
                g = g.to(fs.dtype).contiguous()
            out[ti] = g
        if out and b.flushed:
            raise errors.InvalidExperimentException(
                "a gradient arrived after this batch's gradients were already synchronized (more "
                "backward passes than backward_passes_per_step declared)"
            )
        return out

    def _flush_bucket(self, si: int, bi: int, grads: Dict[int, Any], to_wire: bool, table_filled: bool = False) -> None:
        """Launch the flatten kernel over the gradients collected for this bucket in this pass.  table_filled: the
        bucket's prebuilt dgs_seg table already holds the gradients' addresses (native host loop) and `grads` only
        names the tensors; the caller keeps the gradient tensors alive across this call."""
        fs = self.sets[si]
        b = fs.buckets[bi]
        if not to_wire and fs.acc is None:
            # more passes than configured: grow an accumulator on demand
            fs.acc = torch.zeros(fs.total, dtype=torch.float32, device=self.device)
            fs.acc_ptr = _ptr(fs.acc)
        stream = _stream()
        in_window = fs.in_window
        first_state = in_window[b.tensors[0]]
        if len(grads) == len(b.tensors) and all(in_window[ti] == first_state for ti in b.tensors):
            # the usual case: the whole bucket, all tensors in the same accumulation state -> prebuilt table
            segs = b.segs
            if not table_filled:
                for k, ti in enumerate(b.tensors):
                    segs[k].src = grads[ti].data_ptr()
            mode = (_lib.READ_ACC if first_state else 0) | (_lib.WRITE_WIRE if to_wire else 0)
            self._flatten(fs, b, segs, len(b.tensors), fs.dtype_code, fs.acc_ptr, mode, stream)
        else:
            for read_acc in (False, True):
                idxs = [ti for ti in grads if in_window[ti] == read_acc]
                if not idxs:
                    continue
                segs = (_lib.Seg * len(idxs))()
                for k, ti in enumerate(idxs):
                    t = fs.tensors[ti]
                    seg = segs[k]
                    seg.src, seg.dst_off, seg.numel = grads[ti].data_ptr(), t.off, t.numel
                mode = (_lib.READ_ACC if read_acc else 0) | (_lib.WRITE_WIRE if to_wire else 0)
                self._flatten(fs, b, segs, len(idxs), fs.dtype_code, fs.acc_ptr, mode, stream)
        flags = fs.in_wire if to_wire else fs.in_window
        for ti in grads:
            flags[ti] = True
        # the gradient tensors can be released now: the allocator is stream-ordered
        if to_wire:
            b.flushed = True
            self._reduce_ready(fs)

    def _flatten(self, fs: _FlatSet, b: _Bucket, segs: Any, n: int, grad_code: int, acc_ptr: Any, mode: int, stream: Any) -> None:
        """One flatten launch over tensors of bucket b.  Peer route, wire-writing modes: the values go straight to the
        rank that owns their slice (dgs_flatten_push) -- the reduce-scatter transfer is this kernel's stores."""
        if self.peer and mode & _lib.WRITE_WIRE:
            _lib.calls.dgs_flatten_push(segs, n, grad_code, acc_ptr, fs.peer_wire, fs.wire_code, mode, self.scale * self.unscale,
                                        b.begin, (b.end - b.begin) // self.world, self.world, self.comm.rank, stream)  # type: ignore
        else:
            _lib.calls.dgs_flatten_accumulate(segs, n, grad_code, acc_ptr, fs.wire_ptr, fs.wire_code, mode, self.scale * self.unscale, stream)

    def _reduce_ready(self, fs: Optional[_FlatSet] = None) -> None:
        """Issue allreduces strictly in `_launch_order`, so that every rank issues the same NCCL sequence.
        The order starts as (set, bucket index) and is replaced after the first backward pass by the order
        in which the buckets actually complete (rank 0's, broadcast to all ranks)."""
        order = self._launch_order
        while self._next_launch < len(order):
            si, bi = order[self._next_launch]
            fs = self.sets[si]
            b = fs.buckets[bi]
            if not b.flushed:
                return
            itemsize = fs.wire.element_size()
            base, span = fs.wire.data_ptr() + b.begin * itemsize, b.end - b.begin
            if self.peer:
                # the flatten kernel's stores were the transfer.  Large buckets: wait (on the side stream) until every
                # rank's pushes of THIS bucket have landed and reduce it there, while the compute stream goes on pushing
                # the next buckets; the ranks issue these barriers in the same (agreed) bucket order.  Small sets reduce
                # everything at step time in one launch (finish_window).
                if fs.peer_overlap and not b.reduced:
                    self._peer_reduce_bucket(fs, b)
            elif self.comm is not None and not b.reduced:
                if self.sharded:
                    self.comm.reduce_scatter_raw(base, span // self.world, fs.wire_code)
                else:
                    self.comm.allreduce_raw(base, span, fs.wire_code)
            if self._eager_norm and self.comm is not None and not b.norm_done and not self.peer:
                # behind this bucket's collective, on the communicator's aux stream, while the data is in L2
                if self.sharded:
                    span //= self.world
                    base += self.comm.rank * span * itemsize
                self.comm.l2norm_partial(base, fs.wire_code, span, self._partials.data_ptr() + 4 * b.norm_off)
                b.norm_done = True
            b.reduced = True
            self._next_launch += 1

    def _peer_reduce_bucket(self, fs: _FlatSet, b: _Bucket) -> None:
        assert self.peers is not None and self.comm is not None
        isz = fs.wire.element_size()
        side = self._peer_stream
        if side is not None:
            pushed = torch.cuda.Event()
            pushed.record()  # behind this bucket's push on the compute stream
            side.wait_event(pushed)
            ctx: Any = torch.cuda.stream(side)
        else:
            import contextlib

            ctx = contextlib.nullcontext()  # (CPU emulation: in line)
        with ctx:
            stream = _stream()
            self.peers.barrier(_peer.S_PUSHED, stream)
            _lib.calls.dgs_peer_reduce(_ptr(fs.wire, b.begin * isz), fs.wire_code, (b.end - b.begin) // self.world, self.world,
                                       self.comm.rank, _ptr(self._partials, 4 * b.norm_off), stream)
        b.norm_done = True
        self._peer_side_pending = side is not None

    # ---- step ---------------------------------------------------------------------------------------
    def has_deferred(self) -> bool:
        return any(fs.deferred is not None for fs in self.sets)

    def flush_deferred(self) -> None:
        """Run the flatten that _flush_whole_set left pending: the averaged gradients are needed in the wire buffer
        after all (a clip function that looks at them, a GradScaler, an optimizer whose step stays in torch)."""
        for fs in self.sets:
            if fs.deferred is None:
                continue
            _, read_acc = fs.deferred
            mode = (_lib.READ_ACC if read_acc else 0) | _lib.WRITE_WIRE
            _lib.calls.dgs_flatten_accumulate(fs.segs_all, len(fs.tensors), fs.dtype_code, fs.acc_ptr, fs.wire_ptr, fs.wire_code, mode,
                                              self.scale * self.unscale, _stream())
            fs.deferred = None

    def finish_window(self, allow_direct: bool = False) -> List[List[bool]]:
        """allow_direct: the caller will either call direct_step() or flush_deferred() itself; otherwise pending
        gradients are flattened here, so that everything below finds the wire buffer complete.

        Make every wire slice final (stragglers that never came through a final-pass hook), issue the
        remaining allreduces and make the compute stream wait for them.  Returns, per set, which tensors
        received any gradient in this window."""
        for p in self._frozen_at_wrap:
            if p.grad is not None:
                raise errors.InvalidExperimentException(
                    "a parameter that had requires_grad=False when its optimizer was wrapped has a gradient now: it "
                    "has no place in the optimizer's flat buffers and would never be synchronized or stepped. Unfreeze "
                    "parameters before wrap_optimizer() (a frozen phase can use lr 0 or detach() instead)."
                )
        if sum(len(g["params"]) for g in self.optimizer.param_groups) != self._n_group_params:
            raise errors.InvalidExperimentException(
                "the optimizer's parameters changed after wrap_optimizer() (add_param_group?): its parameters, "
                "gradients and state live in flat buffers laid out when it was wrapped. Create the optimizer with "
                "all its parameter groups before wrapping it (groups can start with lr 0 or frozen parameters)."
            )
        for fs in self.sets:
            base, isz = fs.param.data_ptr(), fs.param.element_size()
            for t in (fs.tensors[0], fs.tensors[-1]):
                if t.param.data_ptr() != base + t.off * isz:
                    raise errors.InvalidExperimentException(
                        "a wrapped optimizer's parameter no longer lives in the optimizer's flat parameter buffer: its "
                        ".data was rebound after wrap_optimizer() (model.to(...), .half(), p.data = ...). The step "
                        "would update memory the model no longer reads. Move or cast the model before wrapping; "
                        "change values in place (p.data.copy_(...), load_state_dict) afterwards."
                    )
        present: List[List[bool]] = []
        self.last_total_norm = None
        if self.has_deferred():
            # gradients produced outside context.backward() since the pending ones were taken
            if _host is not None and hasattr(_host, "any_has_grad"):
                stray = any(_host.any_has_grad(fs.params) for fs in self.sets)
            else:
                stray = any(t.param.grad is not None for fs in self.sets for t in fs.tensors)
            if allow_direct and not stray and all(fs.deferred is not None for fs in self.sets):
                # the usual case on one GPU: every gradient of the window is pending, nothing else to finish
                self._wire_versions = [fs.wire._version for fs in self.sets]
                if self._all_present is None:
                    self._all_present = [[True] * len(fs.tensors) for fs in self.sets]
                return self._all_present
            self.flush_deferred()
        for si, fs in enumerate(self.sets):
            for bi, b in enumerate(fs.buckets):
                grads = self._collect(fs, b)  # gradients produced outside context.backward()
                if grads:
                    self._flush_bucket(si, bi, grads, False)
            have = [fs.in_wire[i] or fs.in_window[i] for i in range(len(fs.tensors))]
            for bi, b in enumerate(fs.buckets):
                if b.flushed:
                    continue
                late = [ti for ti in b.tensors if fs.in_window[ti] and not fs.in_wire[ti]]
                if late:
                    assert fs.acc is not None
                    segs = (_lib.Seg * len(late))()
                    for k, ti in enumerate(late):
                        t = fs.tensors[ti]
                        segs[k].src, segs[k].dst_off, segs[k].numel = fs.acc.data_ptr() + 4 * t.off, t.off, t.numel
                    self._flatten(fs, b, segs, len(late), _lib.F32, None, _lib.WRITE_WIRE, _stream())
                    for ti in late:
                        fs.in_wire[ti] = True
                missing = [ti for ti in b.tensors if not have[ti]]
                if missing:
                    if self.comm is not None:
                        names = len(missing)
                        raise RuntimeError(
                            f"{names} parameter(s) of a wrapped optimizer did not receive a gradient in this "
                            "batch. Distributed training requires every optimizer parameter to take part in "
                            "the loss (torch DDP raises the same complaint for unused parameters)."
                        )
                    for ti in missing:  # absent gradients must not leak stale values into the norm
                        t = fs.tensors[ti]
                        fs.wire[t.off:t.off + t.numel].zero_()
                    self._norm_slots_valid = False
                b.flushed = True
            self._reduce_ready(fs)
            present.append(have)
        if self.peer:
            # every rank's pushes have landed -> each owner sums its rows (rank order) and leaves the per-CTA sums of
            # squares of its reduced slice in the bucket's norm slots
            assert self.peers is not None
            self.peers.check()
            stream = _stream()
            if self._peer_side_pending:
                joined = torch.cuda.Event()
                joined.record(self._peer_stream)
                torch.cuda.current_stream().wait_event(joined)  # the per-bucket reductions issued during backward
                self._peer_side_pending = False
            batched = [fs for fs in self.sets if not fs.peer_overlap]
            if batched:
                self.peers.barrier(_peer.S_PUSHED, stream)
            for fs in batched:
                if len(fs.buckets) <= _lib.PEER_MAX_BUCKETS:
                    if fs.reduce_table is None:
                        bs = sorted(fs.buckets, key=lambda x: x.begin)
                        fs.reduce_table = ((ctypes.c_int64 * len(bs))(*[b.begin for b in bs]),
                                           (ctypes.c_int64 * len(bs))(*[(b.end - b.begin) // self.world for b in bs]),
                                           min(b.norm_off for b in bs), sum(b.norm_n for b in bs))
                    begins, spans, slot0, n_slots = fs.reduce_table
                    _lib.calls.dgs_peer_reduce_multi(fs.wire_ptr, fs.wire_code, begins, spans, len(fs.buckets), self.world, self.comm.rank,  # type: ignore
                                                     _ptr(self._partials, 4 * slot0), n_slots, stream)
                else:
                    isz = fs.wire.element_size()
                    for b in fs.buckets:
                        _lib.calls.dgs_peer_reduce(_ptr(fs.wire, b.begin * isz), fs.wire_code, (b.end - b.begin) // self.world, self.world,
                                                   self.comm.rank, _ptr(self._partials, 4 * b.norm_off), stream)  # type: ignore
                for b in fs.buckets:
                    b.norm_done = True
        elif self.comm is not None:
            self.comm.wait()
        # in-place edits of the gradient views by user code (clip_grad_value_, manual scaling, ...) bump the
        # wire tensors' autograd version; the kernels never do.  total_norm() uses it to notice stale partials.
        self._wire_versions = [fs.wire._version for fs in self.sets]
        return present

    def total_norm(self) -> torch.Tensor:
        """Global L2 norm of the synchronized gradients, left on the device (fp32 scalar).  Call after
        finish_window()."""
        stream = _stream()
        untouched = self._wire_versions == [fs.wire._version for fs in self.sets]
        if self.last_total_norm is not None and untouched:
            return self.last_total_norm  # already measured in this window (AMP's inf check, then the clip function)
        eager_ok = (self.comm is not None and self._eager_norm and self._norm_slots_valid and untouched
                    and all(b.norm_done for fs in self.sets for b in fs.buckets))
        n_bucket_slots = self._partials.numel() - self._bucket_partials_base
        bucket_slots = _ptr(self._partials, 4 * self._bucket_partials_base)
        if self.peer:
            # the owner reduction already wrote the partials of my slices: exchange the ranks' sums (8 bytes each way)
            self.peers.norm_exchange(bucket_slots, n_bucket_slots, _ptr(self._norm), stream)  # type: ignore
        elif self.sharded:
            if not eager_ok:  # pass 1 over this rank's slice of every bucket
                for fs in self.sets:
                    isz = fs.wire.element_size()
                    for b in fs.buckets:
                        span = (b.end - b.begin) // self.world
                        _lib.calls.dgs_l2norm_partial(_ptr(fs.wire, (b.begin + self.comm.rank * span) * isz), fs.wire_code, span,
                                                      _ptr(self._partials, 4 * b.norm_off), stream)
            # sum of squares of my slices -> sum over ranks (one float) -> root
            _lib.calls.dgs_l2norm_final_sumsq(bucket_slots, n_bucket_slots, _ptr(self._norm), stream)
            self.comm.allreduce(self._norm)
            self.comm.wait()
            _lib.calls.dgs_sqrt_inplace(_ptr(self._norm), stream)
        elif eager_ok:
            # pass 1 already ran bucket by bucket behind the allreduces: only the final reduction is left
            _lib.calls.dgs_l2norm_final(bucket_slots, n_bucket_slots, _ptr(self._norm), stream)
        else:
            off = 0
            for fs in self.sets:
                n = _lib.calls.dgs_l2norm_grid(fs.total)
                _lib.calls.dgs_l2norm_partial(_ptr(fs.wire), fs.wire_code, fs.total, _ptr(self._partials, 4 * off), stream)
                off += n
            _lib.calls.dgs_l2norm_final(_ptr(self._partials), off, _ptr(self._norm), stream)
        self._eager_norm, self._eager_unused, self._norm_requested = not self.peer, 0, True
        self.last_total_norm = self._norm[0]
        return self.last_total_norm

    def materialize_grads(self, present: List[List[bool]]) -> None:
        """Expose the synchronized gradients as p.grad views of the wire buffer (user clip functions,
        GradScaler and non-fused optimizers read them)."""
        for fs, have in zip(self.sets, present):
            for t, view, ok in zip(fs.tensors, fs.grad_views, have):
                if not ok:
                    t.param.grad = None
                elif view.dtype == t.param.dtype:
                    t.param.grad = view
                else:
                    t.param.grad = view.to(t.param.dtype)  # fp16-compressed wire, fp32 parameter

    def clear_grads(self) -> None:
        for fs in self.sets:
            for t in fs.tensors:
                t.param.grad = None

    def reset_window(self) -> None:
        if self._eager_norm and not self._norm_requested:
            self._eager_unused += 1
            if self._eager_unused >= 3:
                self._eager_norm = False  # nobody clips by norm any more
        self._norm_slots_valid, self._norm_requested = True, False
        self._next_launch = 0
        for fs in self.sets:
            fs.deferred = None
        if self._pending_order is not None:
            self._launch_order, self._pending_order = self._pending_order, None
        for fs in self.sets:
            for i in range(len(fs.tensors)):
                fs.in_window[i] = fs.in_wire[i] = False
            for b in fs.buckets:
                b.flushed = b.reduced = b.norm_done = False

    # -- fused step -----------------------------------------------------------------------------------
    def can_fuse(self) -> bool:
        return self._fused_kind is not None

    def _state_buffer(self, fs: _FlatSet, key: str) -> torch.Tensor:
        buf = fs.state.get(key)
        if buf is None:
            buf = fs.state[key] = torch.zeros(fs.total, dtype=fs.dtype, device=self.device)
        return buf

    def _adopt_state(self, fs: _FlatSet, key: str) -> Tuple[torch.Tensor, bool]:
        """Flat buffer for optimizer.state[p][key].  State tensors that are not (any more) views of it --
        the user called optimizer.load_state_dict() -- are copied in and replaced by views.  The check
        is by identity on the first and last tensor, so the steady state costs two dict lookups; the
        second result says whether that check passed (every state tensor already was our view)."""
        buf = self._state_buffer(fs, key)
        views = fs.state_views.get(key)
        if views is None:
            views = fs.state_views[key] = [buf[t.off:t.off + t.numel].view(t.param.shape) for t in fs.tensors]
        state = self.optimizer.state
        probes = (fs.tensors[0], fs.tensors[-1])
        if all(state.get(t.param, {}).get(key) is views[t.index] for t in probes):
            return buf, True
        for t in fs.tensors:
            st = state.get(t.param)
            if st is None or st.get(key) is None:
                continue
            if st[key] is not views[t.index]:
                views[t.index].copy_(st[key])
                st[key] = views[t.index]
        return buf, False

    def _ranges(self, fs: _FlatSet, have: Sequence[bool], key_fn: Callable[[_Tensor], Any]) -> List[Tuple[int, int, Any, List[_Tensor]]]:
        """Maximal runs of present tensors with equal key -> (begin, end, key, tensors)."""
        runs: List[Tuple[int, int, Any, List[_Tensor]]] = []
        for t, ok in zip(fs.tensors, have):
            if not ok:
                continue
            k = key_fn(t)
            end = _align(t.off + t.numel)
            if runs and runs[-1][2] == k and runs[-1][1] == t.off:
                b, _, _, ts = runs[-1]
                runs[-1] = (b, end, k, ts + [t])
            else:
                runs.append((t.off, end, k, [t]))
        return runs

    # -- measurement: CUDA events around the optimizer kernel, on the stream it is launched on ----------
    def route_name(self) -> str:
        """Which exchange this engine performs per optimizer step (reported by bench.py)."""
        if self.comm is None:
            return "single-gpu (no exchange)"
        if self.peer:
            return "peer kernels over NVLink: push-flatten (reduce-scatter) + owner reduce + sliced step pushing parameters (all-gather)"
        if self.sharded:
            return "nccl reduce-scatter + sharded step + all-gather"
        return "nccl allreduce" + (" (auto: peer route after two conforming steps)" if self.route == "auto" and self.peers is not None else "")

    def step_kernel_name(self) -> str:
        """The kernel that carries the optimizer step on this engine's current route (reported by bench.py)."""
        kind = {"sgd": "sgd_kernel", "adam": "adam_kernel", "adamw": "adam_kernel"}.get(self._fused_kind or "", "torch optimizer.step()")
        if self._fused_kind is None:
            return kind
        if self.direct and not self.needs_acc:
            return f"dgs::direct_kernel<{self._fused_kind}> (single launch: gradients read in place, flatten + step)"
        return f"dgs::{kind}" + (", PUSH (stores the new parameters to every rank)" if self.peer and not _PUSH_SPLIT else "")

    def step_elements(self) -> int:
        """Elements one optimizer step of this rank covers (all flat sets; 1/N of them with the sharded step)."""
        return sum(fs.total for fs in self.sets) // (self.world if self.sharded else 1)

    def enable_step_timing(self, on: bool = True) -> None:
        """CUDA events recorded inside libdgs immediately around every optimizer-kernel launch."""
        _lib.calls.dgs_timer_enable(1 if on else 0)
        if on:
            self.step_timings_ms()  # drop anything logged earlier

    def step_timings_ms(self) -> List[float]:
        """Durations (ms) of the optimizer-kernel launches since the last call; synchronizes with the last one."""
        buf = (ctypes.c_float * 1024)()
        n = ctypes.c_int(0)
        _lib.calls.dgs_timer_read(buf, 1024, ctypes.byref(n))
        return [buf[i] / 1000.0 for i in range(n.value)]

    def fused_step(self, present: List[List[bool]], clip: Optional[ClipRequest], inv_scale: Optional[torch.Tensor] = None) -> None:
        """clip + optimizer.step() in one kernel per flat set (several if there are > 24 ranges).  `inv_scale`: fp32
        device scalar 1/S when the wire still carries a GradScaler's loss scale; the kernel multiplies it in."""
        assert self._fused_kind is not None
        self._inv_scale = inv_scale
        try:
            self._fused_step(present, clip)
        finally:
            self._inv_scale = None

    def can_direct(self, clip: bool) -> bool:
        """Whether the pending gradients can go through the single-launch step: every flat set deferred, and -- with
        clipping -- one set only (the norm is global; the kernel sees one set)."""
        return bool(self.sets) and all(fs.deferred is not None for fs in self.sets) and (not clip or len(self.sets) == 1)

    def direct_step(self, present: List[List[bool]], max_norm: Optional[float]) -> None:
        """flatten + (norm + clip) + optimizer step of the pending gradients in one kernel per flat set."""
        assert self._fused_kind is not None and self.can_direct(max_norm is not None)
        self._direct_clip = (max_norm is not None, float(max_norm or 0.0))
        try:
            self._fused_step(present, None)
        finally:
            self._direct_clip = None
            for fs in self.sets:
                fs.deferred = None
        if max_norm is not None:
            self.last_total_norm = None

    def _leave_direct(self, norm_ptr: Any, max_norm: float) -> Tuple[Any, float]:
        """The single-launch step does not fit after all (groups with and without momentum, more hyper-parameter rows
        than a launch takes): flatten now and continue on the three-kernel path with the same clipping."""
        self.flush_deferred()
        if self._direct_clip is not None and self._direct_clip[0]:
            self.total_norm()
            return _ptr(self._norm), self._direct_clip[1]
        return norm_ptr, max_norm

    def _direct_launch(self, fs: _FlatSet, entry: str, runs: List[Any], arr: Any, state_ptrs: List[Any]) -> None:
        """runs[k] = (begin, end, key, tensors) is row k of `arr`."""
        if fs.seg_range is None:
            fs.seg_range = (ctypes.c_int32 * len(fs.tensors))()
        shape = tuple(id(run[3]) for run in runs)  # the steady state hands out the cached tensor lists of _group_runs
        if shape != fs.seg_range_for:
            for k, run in enumerate(runs):
                for t in run[3]:
                    fs.seg_range[t.index] = k
            fs.seg_range_for = shape if fs.steady else None
        _, read_acc = fs.deferred  # type: ignore
        clip, max_norm = self._direct_clip  # type: ignore
        n0 = self._partials.numel()
        getattr(_lib.calls, entry)(
            fs.segs_all, fs.seg_range, len(fs.tensors), fs.dtype_code, fs.acc_ptr if read_acc else None, self.scale * self.unscale,
            _ptr(fs.param), *state_ptrs, fs.dtype_code, arr, len(runs), 1 if clip else 0, max_norm,
            _ptr(self._partials), n0, _ptr(self._norm), _stream(),
        )

    def _step_call(self, plain: Any, scaled: Any, *args: Any, push: Any = None) -> None:
        """args = everything up to and including max_norm; the stream goes last.  push = (entry point, destination
        table) on the peer route: the updated parameters are stored to every rank."""
        if push is not None:
            fn, table = push
            self._launch_step(fn, *args, table, self.world, self.comm.rank, _stream())  # type: ignore
        elif self._inv_scale is None:
            self._launch_step(plain, *args, _stream())
        else:
            self._launch_step(scaled, *args, _ptr(self._inv_scale), _stream())

    def _launch_step(self, fn: Any, *args: Any) -> None:
        fn(*args)

    def _push_table(self, fs: _FlatSet, entry: str, lo_bytes: int) -> Any:
        if not self.peer or _PUSH_SPLIT:
            return None
        return getattr(_lib.calls, entry), self.peers.table(f"param{self.sets.index(fs)}", lo_bytes)  # type: ignore

    def _fused_step(self, present: List[List[bool]], clip: Optional[ClipRequest]) -> None:
        norm_ptr, max_norm = None, 0.0
        if clip is not None and clip.max_norm is not None:
            if self.last_total_norm is None:
                self.total_norm()
            norm_ptr, max_norm = _ptr(self._norm), float(clip.max_norm)
        groups = self.optimizer.param_groups
        if self._fused_kind == "sgd":
            self._fused_sgd(present, groups, norm_ptr, max_norm)
        else:
            self._fused_adam(present, groups, norm_ptr, max_norm)
        if self.sharded:
            self._after_sharded_step()
        self.last_total_norm = None
        self.optimizer._opt_called = True  # type: ignore  # what torch LR schedulers look at before warning

    def _step_plan(self, fs: _FlatSet, items: List[Tuple[int, int, Any]]) -> List[Tuple[int, int, List[Tuple[int, int, Any]]]]:
        """Split the (begin, end, key) ranges of one step into kernel launches (lo, hi, ranges): at most
        DGS_MAX_RANGES ranges each, every launch covering the element window [lo, hi) only.  Sharded step: the
        ranges are first cut down to this rank's slice of every bucket, one launch per bucket."""
        plan: List[Tuple[int, int, List[Tuple[int, int, Any]]]] = []
        if not self.sharded:
            for i in range(0, len(items), _lib.MAX_RANGES):
                plan.append((0, fs.total, items[i:i + _lib.MAX_RANGES]))
            return plan
        # one launch over all of this rank's slices (a model has a handful of buckets; a launch per bucket is what
        # bounded small models): the kernel skips tiles outside every range
        rank, world = self.comm.rank, self.world  # type: ignore
        mine: List[Tuple[int, int, Any]] = []
        for b in sorted(fs.buckets, key=lambda x: x.begin):
            span = (b.end - b.begin) // world
            lo, hi = b.begin + rank * span, b.begin + (rank + 1) * span
            mine += [(max(x, lo), min(y, hi), k) for x, y, k in items if x < hi and y > lo]
        if _SLICE_LAUNCH == "per_bucket":  # experiment switch: one launch per slice, no tiles outside the slices
            for b in sorted(fs.buckets, key=lambda x: x.begin):
                span = (b.end - b.begin) // world
                lo, hi = b.begin + rank * span, b.begin + (rank + 1) * span
                part = [r for r in mine if r[0] >= lo and r[1] <= hi]
                for i in range(0, len(part), _lib.MAX_RANGES):
                    chunk = part[i:i + _lib.MAX_RANGES]
                    plan.append((chunk[0][0], chunk[-1][1], chunk))
            return plan
        for i in range(0, len(mine), _lib.MAX_RANGES):
            chunk = mine[i:i + _lib.MAX_RANGES]
            plan.append((chunk[0][0], chunk[-1][1], chunk))
        return plan

    def _after_sharded_step(self) -> None:
        """All-gather the updated parameters, bucket by bucket, in place; the next forward waits for it."""
        assert self.comm is not None
        if self.peer:
            if _PUSH_SPLIT:
                # the step kernels updated this rank's slices in place: copy them to every other rank
                for si, fs in enumerate(self.sets):
                    if fs.scatter_table is None:
                        bs, isz = sorted(fs.buckets, key=lambda x: x.begin), fs.param.element_size()
                        spans = [(b.end - b.begin) // self.world for b in bs]
                        fs.scatter_table = ((ctypes.c_int64 * len(bs))(*[(b.begin + self.comm.rank * sp) * isz for b, sp in zip(bs, spans)]),
                                            (ctypes.c_int64 * len(bs))(*[sp * isz for sp in spans]), len(bs))
                    offs, sizes, n = fs.scatter_table
                    _lib.calls.dgs_peer_scatter(self.peers.table(f"param{si}"), offs, sizes, n, self.world, self.comm.rank, _stream())  # type: ignore
            # the new parameters were stored to every rank: wait until everyone's stores have landed here
            self.peers.barrier(_peer.S_PARAMS, _stream())  # type: ignore
            return
        for fs in self.sets:
            isz = fs.param.element_size()
            for b in sorted(fs.buckets, key=lambda x: x.begin):
                self.comm.allgather_inplace_raw(fs.param.data_ptr() + b.begin * isz, (b.end - b.begin) // self.world * isz)
        self.comm.wait()

    def gather_state(self) -> None:
        """Sharded step only: every rank has advanced the optimizer state of its own slices; bring the other
        ranks' slices in, so that optimizer.state_dict() is complete and identical on every rank (the chief
        writes the checkpoint).  Registered as a state_dict pre-hook of the optimizer."""
        if not self.sharded:
            return
        assert self.comm is not None
        for fs in self.sets:
            for buf in fs.state.values():
                isz = buf.element_size()
                for b in sorted(fs.buckets, key=lambda x: x.begin):
                    self.comm.allgather_inplace_raw(buf.data_ptr() + b.begin * isz, (b.end - b.begin) // self.world * isz)
        self.comm.wait()
        torch.cuda.current_stream().synchronize()

    def _group_runs(self, fs: _FlatSet) -> List[Tuple[int, int, int, List[_Tensor]]]:
        """One (begin, end, group, tensors) per run of adjacent tensors of one param group: the range table
        of a step in which every parameter has a gradient."""
        if fs.group_runs is None:
            fs.group_runs = [(b, e, k, ts) for b, e, k, ts in self._ranges(fs, [True] * len(fs.tensors), lambda t: t.group)]
        return fs.group_runs

    def _fused_sgd(self, present: List[List[bool]], groups: List[Dict[str, Any]], norm_ptr: Any, max_norm: float) -> None:
        for fs, have in zip(self.sets, present):
            any_momentum = any(g["momentum"] != 0 for g in groups)
            buf, intact = self._adopt_state(fs, "momentum_buffer") if any_momentum else (None, True)
            everyone = all(have)
            if everyone and intact and (fs.steady or not any_momentum):
                # steady state: every parameter has a gradient and a momentum buffer that is our view
                runs = [(b, e, (gi, False), ts) for b, e, gi, ts in self._group_runs(fs)]
            else:
                state = self.optimizer.state

                def key(t: _Tensor) -> Any:
                    st = state.get(t.param)
                    return (t.group, st is None or st.get("momentum_buffer") is None)

                runs = self._ranges(fs, have, key)
            psz, wsz = fs.param.element_size(), fs.wire.element_size()
            if fs.deferred is not None:
                kinds = {groups[r[2][0]]["momentum"] != 0 for r in runs}
                if self._direct_clip is not None and everyone and len(kinds) == 1 and len(runs) <= _lib.MAX_RANGES:
                    arr = (_lib.SgdRange * len(runs))()
                    for k, (b, e, (gi, first), _) in enumerate(runs):
                        self._fill_sgd_range(arr[k], b, e, groups[gi], first)
                    self._direct_launch(fs, "dgs_direct_step_sgd", runs, arr, [_ptr(buf) if True in kinds else None])
                    runs_to_launch: List[Any] = []
                else:
                    norm_ptr, max_norm = self._leave_direct(norm_ptr, max_norm)
                    runs_to_launch = runs
            else:
                runs_to_launch = runs
            for with_momentum in (True, False):
                sel = [(r[0], r[1], r[2]) for r in runs_to_launch if (groups[r[2][0]]["momentum"] != 0) == with_momentum]
                for lo, hi, chunk in self._step_plan(fs, sel):
                    arr = (_lib.SgdRange * len(chunk))()
                    for k, (b, e, (gi, first)) in enumerate(chunk):
                        self._fill_sgd_range(arr[k], b - lo, e - lo, groups[gi], first)
                    self._step_call(
                        _lib.calls.dgs_step_sgd, _lib.calls.dgs_step_sgd_scaled,
                        _ptr(fs.param, lo * psz), _ptr(buf, lo * psz) if with_momentum else None,
                        fs.dtype_code, _ptr(fs.wire, lo * wsz), fs.wire_code, hi - lo, arr, len(chunk), norm_ptr, max_norm,
                        push=self._push_table(fs, "dgs_step_sgd_push", lo * psz),
                    )
            if buf is not None:
                views = fs.state_views["momentum_buffer"]
                for b, e, (gi, first), ts in runs:
                    if first and groups[gi]["momentum"] != 0:
                        for t in ts:
                            self.optimizer.state[t.param]["momentum_buffer"] = views[t.index]
                fs.steady = everyone

    @staticmethod
    def _fill_sgd_range(r: Any, begin: int, end: int, g: Dict[str, Any], first: bool) -> None:
        r.begin, r.end = begin, end
        r.lr, r.momentum = float(g["lr"]), float(g["momentum"])
        r.dampening, r.weight_decay = float(g["dampening"]), float(g["weight_decay"])
        r.flags = (
            (_lib.SGD_NESTEROV if g["nesterov"] else 0)
            | (_lib.SGD_FIRST_STEP if first else 0)
            | (_lib.OPT_MAXIMIZE if g.get("maximize") else 0)
        )

    def _fill_adam_range(self, r: Any, begin: int, end: int, g: Dict[str, Any], step: int) -> None:
        b1, b2 = (float(x) for x in g["betas"])
        lr = float(g["lr"])
        r.begin, r.end = begin, end
        r.lr, r.beta1, r.beta2 = lr, b1, b2
        r.eps, r.weight_decay = float(g["eps"]), float(g["weight_decay"])
        r.step_size = lr / (1 - b1 ** step)
        r.bias_correction2_sqrt = (1 - b2 ** step) ** 0.5
        use_decoupled = self._fused_kind == "adamw" or bool(g.get("decoupled_weight_decay"))
        r.flags = (_lib.ADAM_DECOUPLED_WD if use_decoupled else 0) | (_lib.OPT_MAXIMIZE if g.get("maximize") else 0)
        if g["weight_decay"] == 0:
            r.flags |= _lib.ADAM_DECOUPLED_WD  # p *= 1 exactly; skips the L2 add

    def _ensure_steps(self) -> torch.Tensor:
        n = sum(len(fs.tensors) for fs in self.sets)
        if self._steps_host is None:
            self._steps_host = torch.zeros(n, dtype=torch.float32)
            self._step_views = [self._steps_host[i] for i in range(n)]
        return self._steps_host

    def _fused_adam(self, present: List[List[bool]], groups: List[Dict[str, Any]], norm_ptr: Any, max_norm: float) -> None:
        steps = self._ensure_steps()
        base = 0
        decoupled = self._fused_kind == "adamw"
        for fs, have in zip(self.sets, present):
            m, intact_m = self._adopt_state(fs, "exp_avg")
            v, intact_v = self._adopt_state(fs, "exp_avg_sq")
            n_t = len(fs.tensors)
            everyone = all(have)
            if everyone and intact_m and intact_v and fs.steady and fs.uniform_step is not None:
                # steady state: all parameters in lock-step; one host add advances every `step` view
                steps[base:base + n_t] += 1.0
                fs.uniform_step += 1
                runs = [(b, e, (gi, fs.uniform_step), ts) for b, e, gi, ts in self._group_runs(fs)]
            else:
                # per-parameter step counters live in one host tensor; optimizer.state[p]["step"] are views
                state = self.optimizer.state
                for t in fs.tensors:
                    st = state[t.param]
                    cur = st.get("step")
                    mine = self._step_views[base + t.index]
                    if cur is not mine:
                        # no `step` in the optimizer's state: a fresh optimizer, or state that was cleared
                        # (optimizer.state.clear(), load_state_dict of an empty state) -> torch restarts from step 0
                        mine.fill_(float(cur) if cur is not None else 0.0)
                        st["step"] = mine
                    if "exp_avg" not in st:
                        # ... with zero moments: the flat buffers may still hold the old ones
                        fs.state_views["exp_avg"][t.index].zero_()
                        fs.state_views["exp_avg_sq"][t.index].zero_()
                        st["exp_avg"] = fs.state_views["exp_avg"][t.index]
                        st["exp_avg_sq"] = fs.state_views["exp_avg_sq"][t.index]
                steps[base:base + n_t] += torch.tensor(have, dtype=torch.float32)
                counts = [int(c) for c in steps[base:base + n_t].tolist()]
                runs = self._ranges(fs, have, lambda t: (t.group, counts[t.index]))
                fs.steady = everyone
                fs.uniform_step = counts[0] if everyone and len(set(counts)) == 1 else None
            psz, wsz = fs.param.element_size(), fs.wire.element_size()
            if fs.deferred is not None:
                if self._direct_clip is not None and everyone and len(runs) <= _lib.MAX_RANGES:
                    arr = (_lib.AdamRange * len(runs))()
                    for k, (b, e, (gi, step), _) in enumerate(runs):
                        self._fill_adam_range(arr[k], b, e, groups[gi], step)
                    self._direct_launch(fs, "dgs_direct_step_adam", runs, arr, [_ptr(m), _ptr(v)])
                    runs = []
                else:
                    norm_ptr, max_norm = self._leave_direct(norm_ptr, max_norm)
            for lo, hi, chunk in self._step_plan(fs, [(r[0], r[1], r[2]) for r in runs]):
                arr = (_lib.AdamRange * len(chunk))()
                for k, (b, e, (gi, step)) in enumerate(chunk):
                    self._fill_adam_range(arr[k], b - lo, e - lo, groups[gi], step)
                self._step_call(
                    _lib.calls.dgs_step_adam, _lib.calls.dgs_step_adam_scaled,
                    _ptr(fs.param, lo * psz), _ptr(m, lo * psz), _ptr(v, lo * psz), fs.dtype_code,
                    _ptr(fs.wire, lo * wsz), fs.wire_code, hi - lo, arr, len(chunk), norm_ptr, max_norm,
                    push=self._push_table(fs, "dgs_step_adam_push", lo * psz),
                )
            base += n_t
