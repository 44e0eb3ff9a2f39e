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

There is no CPU or torch fallback for these steps.  Optimizers other than SGD / Adam / AdamW keep
their own torch `step()` on gradient views of the wire buffer (still flattened, all-reduced and
clipped by the kernels).
"""
import ctypes
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import torch

from determined_clone_b200 import _lib, errors
from determined_clone_b200.pytorch._comm import Comm

ALIGN_ELEMS = 64  # 128 bytes of bf16, 256 bytes of fp32: every tensor starts vector-aligned
FIRST_BUCKET_BYTES = 1 << 20
BUCKET_BYTES = 25 << 20


def _align(n: int) -> int:
    return (n + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS


def _stream() -> ctypes.c_void_p:
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
    __slots__ = ("begin", "end", "tensors", "arrived", "grads", "flushed", "reduced")

    def __init__(self, begin: int, end: int, tensors: List[int]) -> None:
        self.begin, self.end, self.tensors = begin, end, tensors
        self.arrived = 0
        self.grads: Dict[int, torch.Tensor] = {}
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
        self.next_reduce = 0


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
    ) -> None:
        _lib.load()  # fail now, loudly, if libdgs.so is missing
        if device.type != "cuda":
            raise _lib.DgsError("GradSyncEngine needs a CUDA device: the grad-sync path has no CPU implementation")
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
        self.needs_acc = self.agg * self.bpps > 1
        self.sets: List[_FlatSet] = []
        self._where: Dict[int, Tuple[int, int]] = {}  # id(param) -> (set index, tensor index)
        self._hooks: List[Any] = []
        self._final_pass = False
        self._in_pass = False
        self._norm = torch.zeros(1, dtype=torch.float32, device=device)
        self._partials: torch.Tensor = None  # type: ignore  # scratch of the L2-norm kernels, sized in _build
        self._steps_host: Optional[torch.Tensor] = None  # Adam's per-parameter `step`, one host tensor
        self._step_views: List[torch.Tensor] = []
        self._fused_kind = self._classify(optimizer)
        self.last_total_norm: Optional[torch.Tensor] = None
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
                if not p.requires_grad:
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
            for t in fs.tensors:
                self._hooks.append(t.param.register_post_accumulate_grad_hook(self._on_grad))
        n_partials = sum(_lib.calls.dgs_l2norm_grid(fs.total) for fs in self.sets)
        self._partials = torch.zeros(max(n_partials, 1), dtype=torch.float32, device=self.device)
        if self.comm is not None:
            # what DDP's constructor does: every replica starts from rank 0's parameters
            for fs in self.sets:
                self.comm.broadcast(fs.param, 0)
            self.comm.wait()

    def _cut_buckets(self, fs: _FlatSet) -> None:
        itemsize = torch.empty(0, dtype=fs.wire_dtype).element_size()
        limit = self.first_bucket_bytes
        cur: List[int] = []
        cur_bytes = 0
        end = fs.total
        for t in reversed(fs.tensors):
            cur.append(t.index)
            cur_bytes += t.numel * itemsize
            if cur_bytes >= limit:
                fs.buckets.append(_Bucket(t.off, end, cur))
                end, cur, cur_bytes, limit = t.off, [], 0, self.bucket_bytes
        if cur:
            fs.buckets.append(_Bucket(0, end, cur))
        for bi, b in enumerate(fs.buckets):
            for ti in b.tensors:
                fs.tensors[ti].bucket = bi

    def close(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks = []

    @property
    def parameters(self) -> List[torch.nn.Parameter]:
        return [t.param for fs in self.sets for t in fs.tensors]

    # ---- backward ---------------------------------------------------------------------------------
    def begin_pass(self, final: bool) -> None:
        """Called right before loss.backward().  `final`: this pass completes the window, so finished
        buckets are scaled, cast and all-reduced straight from the hooks."""
        self._in_pass, self._final_pass = True, final
        for fs in self.sets:
            for b in fs.buckets:
                if b.flushed:
                    raise errors.InvalidExperimentException(
                        "context.backward() was called again after the gradients of this batch were "
                        "already synchronized; pass the number of backward calls per batch as "
                        "backward_passes_per_step to wrap_optimizer()"
                    )

    def end_pass(self) -> None:
        """Called right after loss.backward(): flush buckets that only partly filled in this pass."""
        for si, fs in enumerate(self.sets):
            for bi, b in enumerate(fs.buckets):
                if b.grads:
                    self._flush_bucket(si, bi, self._final_pass and b.arrived == len(b.tensors))
        self._in_pass = self._final_pass = False

    def _on_grad(self, p: torch.nn.Parameter) -> None:
        g = p.grad
        if g is None:
            return
        p.grad = None  # the kernel consumes it; autograd must not keep accumulating into it
        si, ti = self._where[id(p)]
        fs = self.sets[si]
        b = fs.buckets[fs.tensors[ti].bucket]
        if b.flushed:
            raise errors.InvalidExperimentException(
                "a gradient arrived after this batch's gradients were already synchronized (more "
                "backward passes than backward_passes_per_step declared)"
            )
        if g.is_sparse:
            raise _lib.DgsError("sparse gradients are not supported by the grad-sync kernels")
        if g.dtype != fs.dtype or not g.is_contiguous():
            g = g.to(fs.dtype).contiguous()
        if ti in b.grads:  # second gradient for the same tensor within one pass: fold on arrival
            b.grads[ti] = b.grads[ti] + g
            return
        b.grads[ti] = g
        b.arrived += 1
        if b.arrived == len(b.tensors):
            self._flush_bucket(si, fs.tensors[ti].bucket, self._final_pass)

    def _flush_bucket(self, si: int, bi: int, to_wire: bool) -> None:
        """Launch the flatten kernel over the gradients this bucket has collected in this pass."""
        fs = self.sets[si]
        b = fs.buckets[bi]
        if not to_wire and fs.acc is None:
            # more passes than configured: grow an accumulator on demand
            fs.acc = torch.zeros(fs.total, dtype=torch.float32, device=self.device)
        for read_acc in (False, True):
            idxs = [ti for ti in b.grads if fs.in_window[ti] == read_acc]
            if not idxs:
                continue
            segs = (_lib.Seg * len(idxs))()
            for k, ti in enumerate(idxs):
                t = fs.tensors[ti]
                segs[k].src, segs[k].dst_off, segs[k].numel = b.grads[ti].data_ptr(), t.off, t.numel
            mode = (_lib.READ_ACC if read_acc else 0) | (_lib.WRITE_WIRE if to_wire else 0)
            _lib.calls.dgs_flatten_accumulate(
                segs, len(idxs), _lib.dtype_code(fs.dtype), _ptr(fs.acc), _ptr(fs.wire),
                _lib.dtype_code(fs.wire_dtype), mode, self.scale, _stream(),
            )
        for ti in b.grads:
            if to_wire:
                fs.in_wire[ti] = True
            else:
                fs.in_window[ti] = True
        b.grads = {}  # stream-ordered allocator: safe to release after the launch
        b.arrived = 0
        if to_wire:
            b.flushed = True
            self._reduce_ready(fs)

    def _reduce_ready(self, fs: _FlatSet) -> None:
        """Issue allreduces strictly in bucket order so that every rank issues the same sequence."""
        while fs.next_reduce < len(fs.buckets) and fs.buckets[fs.next_reduce].flushed:
            b = fs.buckets[fs.next_reduce]
            if self.comm is not None and not b.reduced:
                itemsize = fs.wire.element_size()
                self.comm.allreduce_raw(fs.wire.data_ptr() + b.begin * itemsize, b.end - b.begin, _lib.dtype_code(fs.wire_dtype))
            b.reduced = True
            fs.next_reduce += 1

    # ---- step ---------------------------------------------------------------------------------------
    def finish_window(self) -> List[List[bool]]:
        """Make every wire slice final (stragglers that never came through a final-pass hook), issue the
        remaining allreduces and make the compute stream wait for them.  Returns, per set, which tensors
        received any gradient in this window."""
        present: List[List[bool]] = []
        for si, fs in enumerate(self.sets):
            for bi, b in enumerate(fs.buckets):
                if b.grads:  # gradients produced outside context.backward()
                    self._flush_bucket(si, bi, False)
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
                    _lib.calls.dgs_flatten_accumulate(
                        segs, len(late), _lib.F32, None, _ptr(fs.wire), _lib.dtype_code(fs.wire_dtype),
                        _lib.WRITE_WIRE, self.scale, _stream(),
                    )
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
                b.flushed = True
            self._reduce_ready(fs)
            present.append(have)
        if self.comm is not None:
            self.comm.wait()
        return present

    def total_norm(self) -> torch.Tensor:
        """Global L2 norm of the synchronized gradients, left on the device (fp32 scalar)."""
        off = 0
        for fs in self.sets:
            n = _lib.calls.dgs_l2norm_grid(fs.total)
            _lib.calls.dgs_l2norm_partial(_ptr(fs.wire), _lib.dtype_code(fs.wire_dtype), fs.total, _ptr(self._partials, 4 * off), _stream())
            off += n
        _lib.calls.dgs_l2norm_final(_ptr(self._partials), off, _ptr(self._norm), _stream())
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
        for fs in self.sets:
            fs.next_reduce = 0
            for i in range(len(fs.tensors)):
                fs.in_window[i] = fs.in_wire[i] = False
            for b in fs.buckets:
                b.flushed = b.reduced = False
                b.arrived = 0
                b.grads = {}

    # -- fused step -----------------------------------------------------------------------------------
    def can_fuse(self) -> bool:
        return self._fused_kind is not None

    def _state_buffer(self, fs: _FlatSet, key: str) -> torch.Tensor:
        buf = fs.state.get(key)
        if buf is None:
            buf = fs.state[key] = torch.zeros(fs.total, dtype=fs.dtype, device=self.device)
        return buf

    def _adopt_state(self, fs: _FlatSet, key: str) -> torch.Tensor:
        """Flat buffer for optimizer.state[p][key].  State tensors that are not (any more) views of it --
        the user called optimizer.load_state_dict() -- are copied in and replaced by views.  The check
        is by identity on the first and last tensor, so the steady state costs two dict lookups."""
        buf = self._state_buffer(fs, key)
        views = fs.state_views.get(key)
        if views is None:
            views = fs.state_views[key] = [buf[t.off:t.off + t.numel].view(t.param.shape) for t in fs.tensors]
        state = self.optimizer.state
        probes = (fs.tensors[0], fs.tensors[-1])
        if all(state.get(t.param, {}).get(key) is views[t.index] for t in probes):
            return buf
        for t in fs.tensors:
            st = state.get(t.param)
            if st is None or st.get(key) is None:
                continue
            if st[key] is not views[t.index]:
                views[t.index].copy_(st[key])
                st[key] = views[t.index]
        return buf

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

    def fused_step(self, present: List[List[bool]], clip: Optional[ClipRequest]) -> None:
        """clip + optimizer.step() in one kernel per flat set (several if there are > 24 ranges)."""
        assert self._fused_kind is not None
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
        self.last_total_norm = None

    def _fused_sgd(self, present: List[List[bool]], groups: List[Dict[str, Any]], norm_ptr: Any, max_norm: float) -> None:
        for fs, have in zip(self.sets, present):
            any_momentum = any(g["momentum"] != 0 for g in groups)
            buf = self._adopt_state(fs, "momentum_buffer") if any_momentum else None

            def key(t: _Tensor) -> Any:
                st = self.optimizer.state.get(t.param)
                first = st is None or st.get("momentum_buffer") is None
                return (t.group, first)

            runs = self._ranges(fs, have, key)
            for with_momentum in (True, False):
                sel = [r for r in runs if (groups[r[2][0]]["momentum"] != 0) == with_momentum]
                for i in range(0, len(sel), _lib.MAX_RANGES):
                    chunk = sel[i:i + _lib.MAX_RANGES]
                    arr = (_lib.SgdRange * len(chunk))()
                    for k, (b, e, (gi, first), _) in enumerate(chunk):
                        g = groups[gi]
                        arr[k].begin, arr[k].end = b, e
                        arr[k].lr, arr[k].momentum = float(g["lr"]), float(g["momentum"])
                        arr[k].dampening, arr[k].weight_decay = float(g["dampening"]), float(g["weight_decay"])
                        arr[k].flags = (
                            (_lib.SGD_NESTEROV if g["nesterov"] else 0)
                            | (_lib.SGD_FIRST_STEP if first else 0)
                            | (_lib.OPT_MAXIMIZE if g.get("maximize") else 0)
                        )
                    _lib.calls.dgs_step_sgd(
                        _ptr(fs.param), _ptr(buf) if with_momentum else None, _lib.dtype_code(fs.dtype), _ptr(fs.wire),
                        _lib.dtype_code(fs.wire_dtype), fs.total, arr, len(chunk), norm_ptr, max_norm, _stream(),
                    )
            if buf is not None:
                for b, e, (gi, first), ts in runs:
                    if first and groups[gi]["momentum"] != 0:
                        views = fs.state_views["momentum_buffer"]
                        for t in ts:
                            self.optimizer.state[t.param]["momentum_buffer"] = views[t.index]

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
            m = self._adopt_state(fs, "exp_avg")
            v = self._adopt_state(fs, "exp_avg_sq")
            # per-parameter step counters live in one host tensor; optimizer.state[p]["step"] are views
            for t, ok in zip(fs.tensors, have):
                st = self.optimizer.state[t.param]
                cur = st.get("step")
                mine = self._step_views[base + t.index]
                if cur is not mine:
                    if cur is not None:
                        mine.fill_(float(cur))
                    st["step"] = mine
                    if "exp_avg" not in st:
                        st["exp_avg"] = fs.state_views["exp_avg"][t.index]
                        st["exp_avg_sq"] = fs.state_views["exp_avg_sq"][t.index]
            mask = torch.tensor(have, dtype=torch.float32)
            steps[base:base + len(fs.tensors)] += mask
            counts = steps[base:base + len(fs.tensors)].tolist()

            def key(t: _Tensor) -> Any:
                return (t.group, int(counts[t.index]))

            runs = self._ranges(fs, have, key)
            for i in range(0, len(runs), _lib.MAX_RANGES):
                chunk = runs[i:i + _lib.MAX_RANGES]
                arr = (_lib.AdamRange * len(chunk))()
                for k, (b, e, (gi, step), _) in enumerate(chunk):
                    g = groups[gi]
                    b1, b2 = (float(x) for x in g["betas"])
                    lr = float(g["lr"])
                    arr[k].begin, arr[k].end = b, e
                    arr[k].lr, arr[k].beta1, arr[k].beta2 = lr, b1, b2
                    arr[k].eps, arr[k].weight_decay = float(g["eps"]), float(g["weight_decay"])
                    arr[k].step_size = lr / (1 - b1 ** step)
                    arr[k].bias_correction2_sqrt = (1 - b2 ** step) ** 0.5
                    use_decoupled = decoupled or bool(g.get("decoupled_weight_decay"))
                    arr[k].flags = (_lib.ADAM_DECOUPLED_WD if use_decoupled else 0) | (_lib.OPT_MAXIMIZE if g.get("maximize") else 0)
                    if g["weight_decay"] == 0:
                        arr[k].flags |= _lib.ADAM_DECOUPLED_WD  # p *= 1 exactly; skips the L2 add
                _lib.calls.dgs_step_adam(
                    _ptr(fs.param), _ptr(m), _ptr(v), _lib.dtype_code(fs.dtype), _ptr(fs.wire), _lib.dtype_code(fs.wire_dtype),
                    fs.total, arr, len(chunk), norm_ptr, max_norm, _stream(),
                )
            base += len(fs.tensors)
