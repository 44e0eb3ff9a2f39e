"""PyTorchTrialContext: the object a PyTorchTrial talks to.

Drop-in for the per-step surface of the reference's context
(reference harness/determined/pytorch/_pytorch_context.py): wrap_model :284-311, wrap_optimizer
:313-378, _should_communicate_and_update :711-716, backward :718-816, _average_gradients :818-825,
step_optimizer :827-924, plus the reducer API it inherits (:36, _reducer.py:273-518).  Same
signatures, same exceptions and texts, same gating on aggregation_frequency -- but no torch DDP and no
Horovod underneath: gradients go through GradSyncEngine (libdgs.so kernels + NCCL side stream).

Differences a user can observe, all documented in DESIGN.md:
  * wrap_model returns the model itself on the device (no DistributedDataParallel wrapper, so
    state_dict keys carry no "module." prefix); parameters become views of one flat buffer.
  * between batches p.grad is None (as after zero_grad(set_to_none=True)); inside a clip_grads
    callable p.grad are views of the synchronized, averaged gradient buffer.
  * 16-bit gradients are accumulated in fp32 across micro-batches and rounded once.
"""
import contextlib
import logging
from typing import Any, Callable, Dict, Iterator, List, Optional, Set, Type, Union

import torch
from torch import nn

from determined_clone_b200 import _lib, core, errors, util
from determined_clone_b200.pytorch import _gradsync
from determined_clone_b200.pytorch._comm import Comm
from determined_clone_b200.pytorch._reducer import _PyTorchReducerContext

logger = logging.getLogger("determined.pytorch")


class PyTorchExperimentalContext:
    def __init__(self, parent: Any) -> None:
        self._parent = parent
        self._auto_amp = False
        self._auto_to_device = True
        self._data_repro_checks_disabled = False

    def disable_auto_to_device(self) -> None:
        self._auto_to_device = False

    def disable_dataset_reproducibility_checks(self) -> None:
        """Allow build_*_data_loader to return a plain torch DataLoader (reference _experimental.py:38-61); sharding
        and resuming at the right batch are then the user's responsibility."""
        self._data_repro_checks_disabled = True

    def use_amp(self) -> None:
        """Automatic mixed precision with a default GradScaler (reference _experimental.py:24-36): forward passes
        of models wrapped AFTER this call run under autocast, backward() scales the loss, step_optimizer() unscales,
        checks for inf/nan and steps through the scaler, and the controller updates the scaler at every
        aggregation boundary.  Do not call wrap_scaler() yourself when using this."""
        self._parent.wrap_scaler(torch.amp.GradScaler(self._parent.device.type))
        self._auto_amp = True


def _same_device(a: torch.device, b: torch.device) -> bool:
    """cuda == cuda:0 when one side carries no index."""
    return a.type == b.type and (a.index is None or b.index is None or a.index == b.index)


class _BufferSync:
    """DDP's broadcast_buffers=True: before every training forward, rank 0's module buffers (batch-norm
    running statistics, ...) overwrite every other rank's.  All buffers live in one flat byte tensor so
    that this is one NCCL broadcast."""

    def __init__(self, model: nn.Module, comm: Comm) -> None:
        self.comm = comm
        self.flats: List[torch.Tensor] = []
        items: List[Any] = []
        seen: Set[int] = set()
        for mod in model.modules():
            for name, buf in list(mod._buffers.items()):
                if buf is None or id(buf) in seen or not _same_device(buf.device, comm.device):
                    continue
                seen.add(id(buf))
                items.append((mod, name, buf))
        if items:
            # one byte buffer for buffers of every dtype (16-byte aligned each): ONE broadcast per forward
            offs, total = [], 0
            for _, _, buf in items:
                offs.append(total)
                total += (buf.numel() * buf.element_size() + 15) // 16 * 16
            flat = torch.zeros(total, dtype=torch.uint8, device=items[0][2].device)
            for (mod, name, buf), off in zip(items, offs):
                # alias the storage instead of taking a view: views of one tensor share ONE autograd version
                # counter, and `num_batches_tracked.add_(1)` would then invalidate every running_mean that a
                # batch-norm layer saved for backward
                alias = torch.empty(0, dtype=buf.dtype, device=buf.device).set_(
                    flat.untyped_storage(), off // buf.element_size(), buf.shape, buf.contiguous().stride()
                )
                alias.copy_(buf)
                mod._buffers[name] = alias
            self.flats.append(flat)
        if self.flats:
            self.sync()
            model.register_forward_pre_hook(self._pre_forward)

    def sync(self) -> None:
        for flat in self.flats:
            self.comm.broadcast(flat, 0)
        self.comm.wait()

    def _pre_forward(self, module: nn.Module, args: Any) -> None:
        if module.training and torch.is_grad_enabled():
            self.sync()


class PyTorchTrialContext(_PyTorchReducerContext):
    """Runtime context of one PyTorchTrial.

    Constructor arguments follow the reference (core_context, trial_seed, hparams, slots_per_trial,
    num_gpus, exp_conf, aggregation_frequency, steps_completed, managed_training, debug_enabled,
    enable_tensorboard_logging).  `core_context` only needs a `.distributed` attribute; a bare
    DistributedContext can be given instead through `distributed=`.
    """

    def __init__(
        self,
        core_context: Any = None,
        trial_seed: Optional[int] = None,
        hparams: Optional[Dict] = None,
        slots_per_trial: int = 1,
        num_gpus: int = 1,
        exp_conf: Optional[Dict[str, Any]] = None,
        aggregation_frequency: int = 1,
        steps_completed: int = 0,
        managed_training: bool = True,
        debug_enabled: bool = False,
        enable_tensorboard_logging: bool = False,
        distributed: Optional[core.DistributedContext] = None,
        device: Optional[torch.device] = None,
    ) -> None:
        self._core = core_context
        if distributed is None:
            distributed = getattr(core_context, "distributed", None) or core.DummyDistributedContext()
        self.distributed = distributed
        _PyTorchReducerContext.__init__(self, self.distributed.allgather)
        self._per_slot_batch_size, self._global_batch_size = (
            util.calculate_batch_sizes(hparams=hparams, slots_per_trial=slots_per_trial, trialname="PyTorchTrial")
            if hparams and hparams.get("global_batch_size", None)
            else (None, None)
        )
        self._hparams = hparams
        self._num_gpus = num_gpus
        self._debug_enabled = debug_enabled
        self._exp_conf = exp_conf
        self._trial_seed = trial_seed
        self._steps_completed = steps_completed
        self.device = device if device is not None else self._init_device()
        self._to_device_warned_types: Set[Type] = set()

        self.models: List[nn.Module] = []
        self.optimizers: List[torch.optim.Optimizer] = []
        self.profiler: Any = None
        self.lr_schedulers: List[Any] = []
        self._epoch_len: Optional[int] = None
        self._wrapped_models: Dict[nn.Module, nn.Module] = {}
        self._optimizer_configs: Dict[Any, Any] = {}
        self._main_model = nn.Module()
        self._scaler: Any = None
        self._use_apex = False
        self._loss_ids: Dict[torch.Tensor, int] = {}
        self._last_backward_batch_idx: Optional[int] = None
        self._current_batch_idx: Optional[int] = None

        self.experimental = PyTorchExperimentalContext(self)
        self._managed_training = managed_training
        self._aggregation_frequency = aggregation_frequency
        self._fp16_compression_default = False
        self._average_aggregated_gradients_default = True
        self._is_pre_trainer = False
        self._load_only = False
        self._enable_tensorboard_logging = bool(enable_tensorboard_logging)
        self._stop_requested = False
        self._tensorboard_path: Any = None
        self._tbd_writer: Any = None
        self._determined_profiler: Any = None

        # grad-sync state
        self._engines: Dict[torch.optim.Optimizer, _gradsync.GradSyncEngine] = {}
        self._owned_params: Set[int] = set()
        self._buffer_syncs: List[_BufferSync] = []
        self._backward_calls = 0
        self._backward_calls_batch: Optional[int] = None
        self._comm: Optional[Comm] = None
        # auto-AMP: host copy of the GradScaler's loss scale S, refreshed by an asynchronous read after every
        # scaler.update() (see _loss_scale_on_host)
        self._amp_scale_host: Optional[float] = None
        self._amp_scale_pinned: Optional[torch.Tensor] = None
        self._amp_scale_event: Optional[torch.cuda.Event] = None
        self._amp_scale_used = 1.0

    # ---- device / communicator ----------------------------------------------------------------
    def _init_device(self) -> torch.device:
        if not torch.cuda.is_available():
            # the reference falls back to CPU here (_pytorch_context.py:481-494); the grad-sync path
            # does not exist on CPU, so model code may run but wrap_optimizer() will refuse.
            return torch.device("cpu")
        index = self.distributed.local_rank if self.distributed.size > 1 else torch.cuda.current_device()
        torch.cuda.set_device(index)
        return torch.device("cuda", index)

    def _get_comm(self) -> Optional[Comm]:
        if self.distributed.size < 2:
            return None
        if self._comm is None:
            self._comm = self.distributed.attach_comm(lambda d: Comm.from_distributed(d, self.device))
        return self._comm

    def _set_default_gradient_compression(self, gradient_compression: bool) -> None:
        self._fp16_compression_default = gradient_compression

    def _set_default_average_aggregated_gradients(self, average_aggregated_gradients: bool) -> None:
        self._average_aggregated_gradients_default = average_aggregated_gradients

    # ---- trial-facing helpers -------------------------------------------------------------------
    def get_hparams(self) -> Dict[str, Any]:
        return self._hparams or {}

    def get_hparam(self, name: str) -> Any:
        if name not in self.get_hparams():
            raise ValueError(
                f"Could not find name '{name}' in experiment hyperparameters. Please check your "
                "experiment configuration 'hyperparameters' section."
            )
        return self.get_hparams()[name]

    def get_experiment_config(self) -> Dict[str, Any]:
        if self._exp_conf is None:
            raise ValueError("exp_conf must be passed to the context (or pytorch.init()) to use get_experiment_config()")
        return self._exp_conf

    def get_data_config(self) -> Dict[str, Any]:
        return (self._exp_conf or {}).get("data", {})

    def get_trial_seed(self) -> int:
        if self._trial_seed is None:
            raise ValueError("trial_seed was not set for this context")
        return self._trial_seed

    def get_initial_batch(self) -> int:
        return self._steps_completed

    def get_enable_tensorboard_logging(self) -> bool:
        return self._enable_tensorboard_logging

    def set_enable_tensorboard_logging(self, enable_tensorboard_logging: bool) -> None:
        """Whether the controller also writes the reported metrics as TensorBoard scalars (reference :220-229).
        Uploading the files is the control plane's job and does not exist here; they stay in
        get_tensorboard_path()."""
        if not isinstance(enable_tensorboard_logging, bool):
            raise AssertionError("enable_tensorboard_logging must be a boolean")
        self._enable_tensorboard_logging = enable_tensorboard_logging

    def get_stop_requested(self) -> bool:
        return self._stop_requested

    def set_stop_requested(self, stop_requested: bool) -> None:
        """Ask the training loop to stop after the current batch (reference :200-218; the controller checks it where
        the reference checks preemption, _pytorch_trial.py:521-525)."""
        if not isinstance(stop_requested, bool):
            raise AssertionError("stop_requested must be a boolean")
        if stop_requested:
            logger.info("A trial stoppage has requested. The trial will be stopped at the end of the current step.")
        self._stop_requested = stop_requested

    def get_experiment_id(self) -> int:
        return int(getattr(getattr(self._core, "train", None), "_exp_id", None) or _info_field("experiment_id"))

    def get_trial_id(self) -> int:
        return int(getattr(getattr(self._core, "train", None), "_trial_id", None) or _info_field("trial_id"))

    def get_tensorboard_path(self) -> Any:
        """Directory for TensorBoard files (reference :993-997 asks the core context; without one: a per-trial
        directory under DET_TENSORBOARD_PATH or the system temp dir)."""
        import os
        import pathlib
        import tempfile

        train = getattr(self._core, "train", None)
        if train is not None and hasattr(train, "get_tensorboard_path"):
            return train.get_tensorboard_path()
        if self._tensorboard_path is None:
            base = os.environ.get("DET_TENSORBOARD_PATH") or os.path.join(tempfile.gettempdir(), "determined-tensorboard")
            self._tensorboard_path = pathlib.Path(base) / f"trial-{self.get_trial_id()}" / f"rank-{self.distributed.rank}"
            self._tensorboard_path.mkdir(parents=True, exist_ok=True)
        return self._tensorboard_path

    def get_tensorboard_writer(self) -> Any:
        """A torch.utils.tensorboard.SummaryWriter on get_tensorboard_path(), created on first use (:999-1030)."""
        if self._tbd_writer is None:
            from torch.utils.tensorboard import SummaryWriter

            self._tbd_writer = SummaryWriter(str(self.get_tensorboard_path()))
        return self._tbd_writer

    def _maybe_reset_tbd_writer(self) -> None:
        """Close the writer so that its event file is complete; the next get_tensorboard_writer() opens a new one."""
        if self._tbd_writer is not None:
            self._tbd_writer.close()
            self._tbd_writer = None

    def set_profiler(self, *args: Any, **kwargs: Any) -> None:
        """torch.profiler.profile(*args, **kwargs) with on_trace_ready pointed at the TensorBoard directory; the
        training loop enters it and calls .step() after every batch (reference :424-458, _pytorch_trial.py:891-901)."""
        self.profiler = torch.profiler.profile(
            *args, on_trace_ready=torch.profiler.tensorboard_trace_handler(str(self.get_tensorboard_path())), **kwargs
        )

    def _set_determined_profiler(self, prof: Any) -> None:
        self._determined_profiler = prof

    @contextlib.contextmanager
    def _record_timing(self, metric_name: str, accumulate: bool = False) -> Iterator[None]:
        """Host-clock timing regions of the reference's system profiler (:464-470); a pass-through unless a profiler
        object with record_timing() was attached.  Device-side timing of the optimizer kernels is dgs_timer_*."""
        if not self._determined_profiler:
            yield
            return
        with self._determined_profiler.record_timing(metric_name, accumulate):
            yield

    def _filter_named_parameters(self, optimizer: torch.optim.Optimizer) -> List:
        """(name, parameter) of the wrapped models' parameters that this optimizer owns (:472-479)."""
        mine = {p for group in optimizer.param_groups for p in group.get("params", [])}
        return [(name, p) for name, p in self._main_model.named_parameters() if p in mine]

    def _set_is_pre_trainer(self) -> None:
        self._is_pre_trainer = True

    def configure_apex_amp(self, *args: Any, **kwargs: Any) -> Any:
        # deprecated in the reference (_pytorch_context.py:635-640) and apex is not installed here
        raise errors.InvalidExperimentException(
            "apex.amp is not supported on this path; use torch.amp (autocast + GradScaler via wrap_scaler())"
        )

    def get_per_slot_batch_size(self) -> int:
        if self._per_slot_batch_size is None:
            raise ValueError("global_batch_size not specified in the hyperparameters")
        return self._per_slot_batch_size

    def get_global_batch_size(self) -> int:
        if self._global_batch_size is None:
            raise ValueError("global_batch_size not specified in the hyperparameters")
        return self._global_batch_size

    def current_train_batch(self) -> int:
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        return self._current_batch_idx

    def to_device(self, data: Any) -> Any:
        """Move a (nested) batch to this trial's device with non-blocking copies."""
        if isinstance(data, torch.Tensor):
            return data.to(self.device, non_blocking=True)
        if isinstance(data, dict):
            return {k: self.to_device(v) for k, v in data.items()}
        if isinstance(data, (list, tuple)):
            return type(data)(self.to_device(v) for v in data)
        if type(data) not in self._to_device_warned_types:
            self._to_device_warned_types.add(type(data))
            logger.warning(f"Was not able to move data item of type '{type(data).__name__}' to device.")
        return data

    # ---- wrapping ---------------------------------------------------------------------------------
    def wrap_model(self, model: nn.Module) -> nn.Module:
        """Returns a wrapped model."""
        if self._managed_training:
            if self._use_apex:
                raise errors.InvalidExperimentException("Must call wrap_model() before configure_apex_amp.")
            model = model.to(self.device)
            comm = self._get_comm()
            if comm is not None:
                # stands in for DistributedDataParallel(model): parameters are broadcast when the
                # optimizer is wrapped (they are flattened there), buffers before every forward
                self._buffer_syncs.append(_BufferSync(model, comm))
            self._wrapped_models[model] = model
        model_id = len(self.models)
        self._main_model.__setattr__(f"model_{model_id}", model)
        if self.experimental._auto_amp:
            model = self.autocast_forward_pass(model)
        self.models.append(model)
        return model

    def autocast_forward_pass(self, to_wrap: nn.Module) -> nn.Module:
        """experimental.use_amp(): run the module's forward under autocast and hand fp16 outputs back as fp32, so
        that loss code needs no autocast region of its own (reference _pytorch_context.py:235-282, which builds a
        proxy subclass around the module).  Here the module object stays the one the user made -- attributes,
        state_dict keys and parameter identity are untouched -- and only its bound `forward` is replaced."""
        inner = to_wrap.forward
        device_type = self.device.type
        warned: Set[type] = set()

        def widen(out: Any) -> Any:
            if isinstance(out, torch.Tensor):
                return out.float() if out.dtype == torch.float16 else out
            if isinstance(out, dict):
                plain = {k: widen(v) for k, v in out.items()}
                if type(out) is dict:
                    return plain
                try:  # dict subclasses with keyword constructors (transformers' ModelOutput) keep their type
                    return type(out)(**plain)
                except Exception:
                    return plain
            if isinstance(out, tuple) and hasattr(out, "_fields"):
                return type(out)(*(widen(v) for v in out))
            if isinstance(out, (list, tuple)):
                return type(out)(widen(v) for v in out)
            if type(out) not in warned:
                warned.add(type(out))
                logger.warning(f"Unexpected type '{type(out).__name__}' outputted by model in experimental AMP mode.")
            return out

        def forward(*args: Any, **kwargs: Any) -> Any:
            with torch.autocast(device_type=device_type):
                out = inner(*args, **kwargs)
            return widen(out)

        to_wrap.forward = forward  # type: ignore
        return to_wrap

    def wrap_optimizer(
        self,
        optimizer: torch.optim.Optimizer,
        backward_passes_per_step: int = 1,
        fp16_compression: Optional[bool] = None,
        average_aggregated_gradients: Optional[bool] = None,
    ) -> torch.optim.Optimizer:
        """Returns a wrapped optimizer (the same object: its parameters and state now live in flat
        buffers managed by the grad-sync engine)."""
        if self._managed_training:
            if self._use_apex:
                raise errors.InvalidExperimentException("Must call wrap_optimizer() before configure_apex_amp.")
            if backward_passes_per_step < 1:
                raise errors.InvalidExperimentException(
                    "backward_passes_per_step for local gradient aggregation must be >= 1; "
                    f"got {backward_passes_per_step}."
                )
        if fp16_compression is None:
            fp16_compression = self._fp16_compression_default
        if average_aggregated_gradients is None:
            average_aggregated_gradients = self._average_aggregated_gradients_default
        self._optimizer_configs[optimizer] = {"average_aggregated_gradients": average_aggregated_gradients}
        if optimizer in self._engines:
            raise errors.InvalidExperimentException("wrap_optimizer() was called twice on the same optimizer")
        mine = {id(p) for g in optimizer.param_groups for p in g["params"]}
        if mine & self._owned_params:
            raise errors.InvalidExperimentException(
                "a parameter belongs to two wrapped optimizers; each parameter's gradient and state can live in "
                "only one optimizer's flat buffers"
            )
        self._owned_params |= mine
        if self._load_only:  # CheckpointLoadContext: the optimizer is registered so that its state can be restored
            self.optimizers.append(optimizer)
            return optimizer
        # managed or not (Trainer / Core API use): the same engine carries the step
        self._engines[optimizer] = _gradsync.GradSyncEngine(
            optimizer,
            device=self.device,
            comm=self._get_comm(),
            aggregation_frequency=self._aggregation_frequency,
            average_aggregated_gradients=bool(average_aggregated_gradients),
            backward_passes_per_step=backward_passes_per_step,
            fp16_compression=bool(fp16_compression) and self.distributed.size > 1,
            sharded_step=((self._exp_conf or {}).get("optimizations", {}) or {}).get("sharded_step"),
            route=((self._exp_conf or {}).get("optimizations", {}) or {}).get("grad_sync_route"),
        )
        self.optimizers.append(optimizer)
        return optimizer

    def wrap_lr_scheduler(self, lr_scheduler: Any, step_mode: Any = None, frequency: int = 1) -> Any:
        from determined_clone_b200.pytorch._lr_scheduler import LRScheduler

        opt = getattr(lr_scheduler, "optimizer", None)
        if opt is not None and opt not in self.optimizers:
            raise errors.InvalidExperimentException("Must use an optimizer that is returned by wrap_optimizer()")
        wrapped = LRScheduler(lr_scheduler, step_mode if step_mode is not None else LRScheduler.StepMode.STEP_EVERY_BATCH, frequency)
        self.lr_schedulers.append(wrapped)
        return lr_scheduler

    def wrap_scaler(self, scaler: Any) -> Any:
        self._scaler = scaler
        return scaler

    # ---- gating -------------------------------------------------------------------------------------
    @contextlib.contextmanager
    def _no_sync(self) -> Iterator[None]:
        # DDP's no_sync has nothing to switch off here: whether a backward pass communicates is
        # decided per pass in backward().
        yield

    def _should_communicate_and_update(self) -> bool:
        if not self._managed_training:
            return True
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        return (self._current_batch_idx + 1) % self._aggregation_frequency == 0

    # ---- backward ---------------------------------------------------------------------------------
    def backward(
        self,
        loss: torch.Tensor,
        gradient: Optional[torch.Tensor] = None,
        retain_graph: bool = False,
        create_graph: bool = False,
    ) -> None:
        """Compute the gradient of `loss` w.r.t. graph leaves (arguments as torch.Tensor.backward).

        On micro-batches that are not an aggregation boundary the gradients are only accumulated into
        the engines' fp32 buckets.  On the boundary pass each bucket is scaled by 1/(N*A), cast and
        all-reduced as soon as it is complete, while autograd keeps running.
        """
        if self._scaler and self.experimental._auto_amp:
            loss = self._scaler.scale(loss)
        if self._backward_calls_batch != self._current_batch_idx:
            self._backward_calls_batch, self._backward_calls = self._current_batch_idx, 0
        self._backward_calls += 1
        boundary = self._should_communicate_and_update() if self._engines else False
        unscale = 1.0 / self._loss_scale_on_host() if self._auto_amp_active() else 1.0
        # An optimizer that has already stepped in THIS batch sits the rest of the batch out: gradients that later
        # passes leave on its parameters (a generator loss flowing through the discriminator) stay in p.grad, as in
        # the reference, where the trial may clear them or let the next batch accumulate onto them.
        batch = self._current_batch_idx
        active = [eng for eng in self._engines.values() if batch is None or eng.stepped_in_batch != batch]
        for eng in active:
            # auto-AMP: 1/S rides on the 1/(N*A) factor of the flatten kernel, so the wire holds true gradients
            # and GradScaler.unscale_'s read+write pass over every gradient never runs
            eng.unscale = unscale if self._prefolds(eng) else 1.0
            eng.begin_pass(final=boundary and self._backward_calls >= eng.bpps)
        try:
            loss.backward(gradient=gradient, retain_graph=retain_graph, create_graph=create_graph)  # type: ignore
        finally:
            for eng in active:
                eng.end_pass()

    @staticmethod
    def _average_gradients(parameters: Any, divisor: int) -> None:
        """Reference semantics on explicit .grad tensors (kept for callers that use it directly; the
        engine folds this division into the scale factor of its flatten kernel)."""
        if divisor == 1:
            return
        divisor_value = float(divisor)
        for p in parameters:
            if p.grad is not None:
                p.grad.data.div_(divisor_value)

    # ---- AMP ------------------------------------------------------------------------------------------
    def _auto_amp_active(self) -> bool:
        return bool(self._scaler) and self.experimental._auto_amp and self._scaler.is_enabled()

    def _prefolds(self, eng: Any) -> bool:
        """Whether this engine's gradients are unscaled by the flatten kernel (auto-AMP + a step we fuse)."""
        return (self._auto_amp_active() and eng.can_fuse() and not eng.sharded
                and all(fs.wire_dtype == fs.dtype for fs in eng.sets))

    def _loss_scale_on_host(self) -> float:
        """S of the auto-AMP scaler, without a device round trip in steady state: S only changes in
        scaler.update(), after which _scaler_updated() starts an asynchronous read into pinned memory; by the next
        boundary backward pass (a whole forward/backward later) that read is long complete.  step_optimizer()
        cross-checks the value against the device copy in the read-back it does anyway."""
        if self._amp_scale_event is not None:
            self._amp_scale_event.synchronize()
            self._amp_scale_host = float(self._amp_scale_pinned[0])  # type: ignore
            self._amp_scale_event = None
        elif self._amp_scale_host is None:
            self._amp_scale_host = float(self._scaler.get_scale())  # first window: one blocking read
        self._amp_scale_used = self._amp_scale_host
        return self._amp_scale_host

    def _scaler_updated(self) -> None:
        """The controller calls this right after scaler.update() (reference _pytorch_trial.py:903-910)."""
        scale = getattr(self._scaler, "_scale", None)
        if not self._auto_amp_active() or scale is None or not scale.is_cuda:
            self._amp_scale_host = None
            return
        if self._amp_scale_pinned is None:
            self._amp_scale_pinned = torch.zeros(1, dtype=torch.float32).pin_memory()
        self._amp_scale_pinned.copy_(scale.reshape(1), non_blocking=True)
        self._amp_scale_event = torch.cuda.Event()
        self._amp_scale_event.record()

    def _amp_fused_step(self, eng: Any, optimizer: Any, parameters: List[Any], present: Any, clip_grads: Any,
                        scaler: Any, auto_zero_grads: bool) -> None:
        """step_optimizer with a GradScaler when the optimizer step is ours (reference :899-921: unscale_, the
        user's clip function, scaler.step -> optimizer.step unless an inf/nan was found).  Same observable
        protocol as torch's -- per-optimizer stage, found_inf recorded for scaler.update(), ONE host read per
        step to decide whether to step (torch: found_inf.item() in GradScaler._maybe_opt_step) -- but the
        unscale never runs as a pass of its own and clip + step stay one kernel:

        * auto-AMP (experimental.use_amp): the flatten kernel already divided by S; the inf/nan check is the
          L2-norm pass (non-finite norm <=> some element is inf/nan), which is also the clip norm.
        * a scaler passed to step_optimizer(): gradients are still scaled.  If the user's clip function called
          scaler.unscale_() itself, torch has unscaled the views in place and recorded found_inf; otherwise the
          norm pass is the check and the step kernel multiplies by 1/S (dgs_step_*_scaled)."""
        import math

        from torch.amp.grad_scaler import OptState

        scaler._check_scale_growth_tracker("step")
        st = scaler._per_optimizer_states[id(optimizer)]
        if st["stage"] is OptState.STEPPED:
            raise RuntimeError("step() has already been called since the last update().")
        prefolded = scaler is self._scaler and self._prefolds(eng)
        clip_req: Optional[_gradsync.ClipRequest] = None
        inv_scale: Optional[torch.Tensor] = None
        found: Optional[bool] = None
        materialized = False

        if prefolded:
            norm_val, s_dev = torch.stack([eng.total_norm(), scaler._scale.reshape(())]).tolist()
            if s_dev != self._amp_scale_used:
                # S changed behind the controller's back (a manual update() or load_state_dict()): rescale once
                fix = self._amp_scale_used / s_dev
                for fs in eng.sets:
                    fs.wire.mul_(fix)
                eng._norm.mul_(fix)
                eng._wire_versions = [fs.wire._version for fs in eng.sets]
                norm_val *= fix
                self._amp_scale_host = None
            found = not math.isfinite(norm_val)
            st["found_inf_per_device"] = {self.device: torch.full((), float(found), dtype=torch.float32, device=self.device)}
            st["stage"] = OptState.UNSCALED  # what the user's function would see after the reference's unscale_

        if clip_grads is not None or not auto_zero_grads:
            eng.materialize_grads(present)
            materialized = True
        if clip_grads is not None:
            if not prefolded and self._scaler and self.experimental._auto_amp:
                self._scaler.unscale_(optimizer)
            clip_req = _run_clip(eng, clip_grads, parameters, defer=auto_zero_grads and _is_plain_norm_clip(clip_grads))
            _fold_back_replaced_grads(eng, present)

        if found is None and st["stage"] is OptState.UNSCALED:
            # torch unscaled the gradient views in place and checked them
            found = bool(sum(v.item() for v in st["found_inf_per_device"].values()))
        elif found is None:
            if clip_req is not None and clip_req.max_norm is not None:
                # clip_grad_norm_ on still-scaled gradients: honoured literally (max_norm against the scaled norm)
                _apply_clip_now(eng, clip_req.max_norm)
                clip_req = None
                norm = eng._norm[0]
            else:
                norm = eng.total_norm()
            found = not math.isfinite(float(norm))
            st["found_inf_per_device"] = {self.device: torch.full((), float(found), dtype=torch.float32, device=self.device)}
            inv_scale = scaler._scale.double().reciprocal().float()

        if not found:
            eng.fused_step(present, clip_req, inv_scale=inv_scale)
        else:
            # skipped.  A whole-buffer clip with a non-finite coefficient has also turned the alignment gaps between
            # tensors into NaN; the flatten kernel never rewrites those and the norm pass reads them
            for fs in eng.sets:
                fs.wire.zero_()
        st["stage"] = OptState.STEPPED
        if auto_zero_grads and materialized:
            eng.clear_grads()
        eng.reset_window()
        eng.stepped_in_batch = self._current_batch_idx

    # ---- optimizer step -----------------------------------------------------------------------------
    def step_optimizer(
        self,
        optimizer: torch.optim.Optimizer,
        clip_grads: Optional[Callable[[Iterator], None]] = None,
        auto_zero_grads: bool = True,
        scaler: Optional[Any] = None,
    ) -> None:
        """Perform a single optimization step: synchronize, average, clip, step, zero.

        A no-op on batches that are not an aggregation boundary.  `clip_grads` receives the list of
        parameters; if it calls torch.nn.utils.clip_grad_norm_ on them the norm comes from the L2-norm
        kernels and the clipping is fused into the optimizer kernel; anything else it does to the
        p.grad views is honoured as is.
        """
        if self._aggregation_frequency > 1 and not auto_zero_grads:
            raise errors.InvalidExperimentException(
                "if optimizations.aggregation_frequency is larger than 1, "
                "auto_zero_grads must be set to true. ",
            )
        if not self._should_communicate_and_update():
            return
        eng = self._engines.get(optimizer)
        if eng is None:
            if self._load_only:
                raise errors.InvalidExperimentException("a CheckpointLoadContext only loads trials; use pytorch.init() to train")
            raise errors.InvalidExperimentException("step_optimizer() needs an optimizer returned by wrap_optimizer()")

        present = eng.finish_window(allow_direct=True)
        parameters = [p for group in optimizer.param_groups for p in group.get("params", [])]
        if scaler is None and self.experimental._auto_amp:
            scaler = self._scaler
        if scaler is not None and not scaler.is_enabled():
            scaler = None  # a disabled GradScaler is a pass-through in torch too
        # what the sliced routes (every rank holds and steps 1/N of each bucket) can honour; see GradSyncEngine.route
        sliceable_call = scaler is None and auto_zero_grads and eng.can_fuse()
        if eng.has_deferred():
            # one GPU: the gradients of this window are still where autograd left them.  If nothing needs them in memory
            # (our optimizer kernel steps, no GradScaler, a clip function that provably never looks) the whole tail --
            # average, norm, clip, step -- is one kernel over them; otherwise they are flattened now, as before.
            direct, max_norm = sliceable_call, None
            if direct and clip_grads is not None:
                max_norm = _plain_clip_max_norm(eng, clip_grads, parameters) if _is_plain_norm_clip(clip_grads) else None
                direct = max_norm is not None
            if direct and eng.can_direct(max_norm is not None):
                eng.direct_step(present, max_norm)
                eng.reset_window()
                eng.stepped_in_batch = self._current_batch_idx
                return
            eng.flush_deferred()
        if eng.sharded:
            if eng.route == "auto" and not sliceable_call:
                eng.demote(present)  # back to full gradients on every rank, for good; this step continues below
            else:
                return self._sharded_step(eng, optimizer, parameters, present, clip_grads, scaler)
        aliased = all(fs.wire_dtype == fs.dtype for fs in eng.sets)
        if scaler is not None and eng.can_fuse() and (aliased or clip_grads is None):
            eng.note_step(False)
            return self._amp_fused_step(eng, optimizer, parameters, present, clip_grads, scaler, auto_zero_grads)
        fused = eng.can_fuse() and scaler is None and (aliased or clip_grads is None)

        clip_req: Optional[_gradsync.ClipRequest] = None
        materialized = False
        pure_clip = clip_grads is None
        if clip_grads is not None or not fused or not auto_zero_grads:
            eng.materialize_grads(present)
            materialized = True
        if clip_grads is not None:
            if self._scaler and self.experimental._auto_amp:
                self._scaler.unscale_(optimizer)
            versions = [fs.wire._version for fs in eng.sets]
            # The clip is folded into the optimizer kernel only for a function that provably does nothing else (and when
            # the gradients are dropped afterwards): anything the function did to the views AFTER clip_grad_norm_, or
            # a user who keeps p.grad, must see clipped values, as in torch -- then the clip is one more pass, now.
            clip_req = _run_clip(eng, clip_grads, parameters, defer=fused and aliased and auto_zero_grads and _is_plain_norm_clip(clip_grads))
            # exactly clip_grad_norm_ over the optimizer's parameters, nothing edited in place, nothing replaced
            pure_clip = (clip_req.max_norm is not None and versions == [fs.wire._version for fs in eng.sets]
                         and all(t.param.grad is v for fs, have in zip(eng.sets, present)
                                 for t, v, ok in zip(fs.tensors, fs.grad_views, have) if ok))
            if aliased:
                _fold_back_replaced_grads(eng, present)

        if fused:
            eng.fused_step(present, clip_req)
        elif scaler:
            scaler.step(optimizer)
        else:
            optimizer.step()

        if auto_zero_grads:
            if materialized:
                eng.clear_grads()
        eng.reset_window()
        eng.stepped_in_batch = self._current_batch_idx
        # the sliced routes cannot show a clip function the whole gradient: only a function that provably never looks
        # (none at all, or nothing but clip_grad_norm_) lets route "auto" move there
        eng.note_step(sliceable_call and fused and pure_clip and (clip_grads is None or _is_plain_norm_clip(clip_grads)))

    def _sharded_step(self, eng: Any, optimizer: Any, parameters: List[Any], present: Any, clip_grads: Any, scaler: Any) -> None:
        """step_optimizer when the engine reduce-scatters (opt-in sharded step): each rank holds only its slice
        of the averaged gradients, so the only clip function that can be honoured is the one the kernels
        implement -- torch.nn.utils.clip_grad_norm_ over all of the optimizer's parameters."""
        if scaler:
            raise errors.InvalidExperimentException(
                "sharded_step cannot be combined with a GradScaler (set optimizations.grad_sync_route: allreduce)")
        clip_req = None
        if clip_grads is not None:
            eng.materialize_grads(present)
            versions = [fs.wire._version for fs in eng.sets]
            clip_req = _run_clip(eng, clip_grads, parameters, defer=True)
            eng.clear_grads()
            if versions != [fs.wire._version for fs in eng.sets]:  # the function edited gradients in place
                raise errors.InvalidExperimentException(
                    "with optimizations.sharded_step (or the peer route that grad_sync_route: auto moves to after two "
                    "optimizer steps whose clip_grads only called clip_grad_norm_) the clip_grads function may only call "
                    "torch.nn.utils.clip_grad_norm_(params, max_norm) on the parameters it is given: every rank "
                    "holds just its own slice of the averaged gradients. Set optimizations.grad_sync_route: allreduce "
                    "for a clip function that edits gradients in place on some steps only."
                )
        eng.fused_step(present, clip_req)
        eng.reset_window()
        eng.stepped_in_batch = self._current_batch_idx

    # ---- epoch helpers the controller maintains ---------------------------------------------------
    def is_epoch_start(self) -> bool:
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        if self._epoch_len is None:
            raise errors.InternalException("Training DataLoader uninitialized.")
        return self._current_batch_idx % self._epoch_len == 0

    def is_epoch_end(self) -> bool:
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        if self._epoch_len is None:
            raise errors.InternalException("Training DataLoader uninitialized.")
        return self._current_batch_idx % self._epoch_len == self._epoch_len - 1

    def current_train_epoch(self) -> int:
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        if self._epoch_len is None:
            raise errors.InternalException("Training DataLoader uninitialized.")
        return self._current_batch_idx // self._epoch_len

    def close(self) -> None:
        for eng in self._engines.values():
            eng.close()


def _info_field(name: str) -> int:
    """experiment_id / trial_id from the cluster info the launcher exported (DET_* environment), else 0."""
    from determined_clone_b200 import _info

    info = _info.get_cluster_info()
    trial = getattr(info, "trial", None) if info is not None else None
    return int(getattr(trial, name, 0) or 0)


_PLAIN_CLIP_NAMES = {"torch", "nn", "utils", "clip_grad", "clip_grad_norm_"}


def _is_plain_norm_clip(fn: Any) -> bool:
    """True only if `fn` provably does nothing but call torch.nn.utils.clip_grad_norm_ on its argument -- the one
    clip function the sliced routes (each rank holds 1/N of the averaged gradients) can serve, because it never gets
    to READ a gradient.  Decided statically and conservatively: a functools.partial of clip_grad_norm_, or a plain
    function / lambda / method whose code names nothing outside {torch, nn, utils, clip_grad, clip_grad_norm_}, has no
    nested code and closes over numbers only (the reference's own example is such a lambda, reference
    harness/determined/pytorch/_pytorch_context.py:851-853).  Anything else -- a callable object, a function that
    logs gradient statistics -- answers False and keeps full gradients on every rank."""
    import functools
    import types

    real = torch.nn.utils.clip_grad_norm_
    if isinstance(fn, functools.partial):
        return fn.func is real and not fn.args and all(isinstance(v, (int, float, bool, type(None))) for v in fn.keywords.values())
    fn = getattr(fn, "__func__", fn)  # bound method -> function
    code = getattr(fn, "__code__", None)
    if not isinstance(fn, types.FunctionType) or code is None:
        return False
    if any(isinstance(c, types.CodeType) for c in code.co_consts) or not set(code.co_names) <= _PLAIN_CLIP_NAMES:
        return False
    if "clip_grad_norm_" not in code.co_names and "clip_grad_norm_" not in code.co_freevars:
        return False
    import dis

    if any(ins.opname in ("FOR_ITER", "GET_ITER", "JUMP_BACKWARD", "YIELD_VALUE") for ins in dis.get_instructions(code)):
        return False  # a loop (comprehensions are inlined since 3.12): per-parameter calls are not the one global clip
    expected = {"torch": torch, "nn": torch.nn, "utils": torch.nn.utils, "clip_grad": torch.nn.utils.clip_grad, "clip_grad_norm_": real}
    # what the function closes over: numbers (the max_norm), or the very objects the allowed names stand for
    # (`from torch.nn.utils import clip_grad_norm_` inside an enclosing function)
    for name, cell in zip(code.co_freevars, fn.__closure__ or ()):
        try:
            value = cell.cell_contents
        except ValueError:
            return False
        if isinstance(value, bool) or not (isinstance(value, (int, float)) or (name in expected and value is expected[name])):
            return False
    if "clip_grad_norm_" in code.co_freevars:
        return True  # reached through the closure, checked above; the remaining names are attribute walks
    g = fn.__globals__
    root = code.co_names[0]  # the first name loaded is a global (or builtin) lookup; the rest are attribute walks from it
    return g.get(root) is expected[root]


def _plain_clip_max_norm(eng: _gradsync.GradSyncEngine, clip_grads: Callable, parameters: List[Any]) -> Optional[float]:
    """max_norm of a clip function that _is_plain_norm_clip vouched for, found by running it with torch's
    clip_grad_norm_ internals replaced by recorders (the parameters carry no .grad at this point, so it has nothing
    to touch).  None if it asked for anything but one plain L2 clip (another norm_type, error_if_nonfinite, two
    calls): the caller then takes the general path."""
    from torch.nn.utils import clip_grad as tcg

    calls: List[Any] = []
    real_norm, real_clip = tcg._get_total_norm, tcg._clip_grads_with_norm_

    def total_norm(tensors: Any, norm_type: float = 2.0, error_if_nonfinite: bool = False, foreach: Any = None) -> torch.Tensor:
        calls.append(("norm", float(norm_type), bool(error_if_nonfinite)))
        return eng._norm[0]  # the kernel writes the norm here; a plain clip function discards the value anyway

    def clip_with_norm(params: Any, max_norm: float, total: torch.Tensor, foreach: Any = None) -> None:
        calls.append(("clip", float(max_norm)))

    tcg._get_total_norm, tcg._clip_grads_with_norm_ = total_norm, clip_with_norm  # type: ignore
    try:
        clip_grads(parameters)
    finally:
        tcg._get_total_norm, tcg._clip_grads_with_norm_ = real_norm, real_clip
    if len(calls) == 2 and calls[0] == ("norm", 2.0, False) and calls[1][0] == "clip":
        return calls[1][1]
    return None


def _fold_back_replaced_grads(eng: _gradsync.GradSyncEngine, present: List[List[bool]]) -> None:
    """A clip function may assign a new tensor to p.grad instead of modifying it in place; copy such
    tensors back into the wire buffer the optimizer kernel reads."""
    for fs, have in zip(eng.sets, present):
        for t, view, ok in zip(fs.tensors, fs.grad_views, have):
            g = t.param.grad
            if ok and g is not view and g is not None:
                view.copy_(g)
                t.param.grad = view


def _apply_clip_now(eng: _gradsync.GradSyncEngine, max_norm: float) -> None:
    """g *= min(max_norm / (norm + 1e-6), 1) over the wire, with the norm the engine measured last."""
    for fs in eng.sets:
        _lib.calls.dgs_clip_scale(
            _gradsync._ptr(fs.wire), _lib.dtype_code(fs.wire_dtype), fs.total, _gradsync._ptr(eng._norm),
            float(max_norm), _gradsync._stream(),
        )
    eng.last_total_norm = None  # the buffer changed under the cached value


def _run_clip(eng: _gradsync.GradSyncEngine, clip_grads: Callable, parameters: List[Any], defer: bool) -> _gradsync.ClipRequest:
    """Run the user's clip function with torch's clip_grad_norm_ internals redirected to the kernels.

    torch.nn.utils.clip_grad_norm_ computes `_get_total_norm(grads, ...)` then
    `_clip_grads_with_norm_(parameters, max_norm, total_norm, ...)`, both looked up in its module at
    call time -- so redirecting those two names catches every way a user may have imported
    clip_grad_norm_.  Calls that are not "L2 norm over exactly this optimizer's gradients" fall
    through to torch untouched.
    """
    from torch.nn.utils import clip_grad as tcg

    req = _gradsync.ClipRequest()
    ours = [v for fs in eng.sets for t, v in zip(fs.tensors, fs.grad_views) if t.param.grad is v]
    aliased = len(ours) > 0
    real_norm, real_clip = tcg._get_total_norm, tcg._clip_grads_with_norm_
    issued: List[torch.Tensor] = []

    def total_norm(tensors: Any, norm_type: float = 2.0, error_if_nonfinite: bool = False, foreach: Any = None) -> torch.Tensor:
        ts = [tensors] if isinstance(tensors, torch.Tensor) else list(tensors)
        if aliased and float(norm_type) == 2.0 and not error_if_nonfinite and len(ts) == len(ours) and all(a is b for a, b in zip(ts, ours)):
            out = eng.total_norm().clone()  # the caller may keep it; the engine's scalar is reused every step
            issued.append(out)
            return out
        return real_norm(ts, norm_type, error_if_nonfinite, foreach)

    def clip_with_norm(params: Any, max_norm: float, total: torch.Tensor, foreach: Any = None) -> None:
        if issued and total is issued[-1] and req.max_norm is None:
            if defer:
                req.max_norm = float(max_norm)  # fused into the optimizer kernel
            else:
                _apply_clip_now(eng, max_norm)
            return
        real_clip(params, max_norm, total, foreach)

    tcg._get_total_norm, tcg._clip_grads_with_norm_ = total_norm, clip_with_norm  # type: ignore
    try:
        clip_grads(parameters)
    finally:
        tcg._get_total_norm, tcg._clip_grads_with_norm_ = real_norm, real_clip
    return req
